// common.cuh — small device helpers shared by the sm_100a kernels of libmvl_b200.
// Everything here is written for sm_100a only (packed fp32 FFMA2, cp.async, TMA bulk copies).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mvl {

typedef unsigned long long u64;

// ---------------------------------------------------------------- packed fp32 pairs (FFMA2)
__device__ __forceinline__ u64 pack2(float lo, float hi) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// d = a * b + c on two fp32 lanes with one instruction (SASS: FFMA2).
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) {
  u64 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) {
  u64 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float hsum2(u64 v) {
  float lo, hi;
  unpack2(v, lo, hi);
  return lo + hi;
}

// ---------------------------------------------------------------- shared-memory access by address
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(a));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w));
}
__device__ __forceinline__ void sts64(uint32_t a, uint2 v) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(v.x), "r"(v.y));
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v));
}

// base + row * pitch as one IMAD.WIDE.U32 (row and pitch are 32-bit, the sum is a 64-bit address).
__device__ __forceinline__ const char* row_ptr(const char* base, uint32_t row, uint32_t pitch) {
  u64 r;
  asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(row), "r"(pitch), "l"(base));
  return reinterpret_cast<const char*>(r);
}
// a ^ imm, volatile so the compiler recomputes it where it is used instead of hoisting a family of
// loop-invariant swizzled addresses into (spilled) registers.
__device__ __forceinline__ uint32_t xor_imm(uint32_t a, uint32_t imm) {
  uint32_t r;
  asm volatile("xor.b32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(imm));
  return r;
}

// ---------------------------------------------------------------- cp.async (SASS: LDGSTS)
// 16-byte global->shared copy that bypasses L1; src_bytes = 0 zero-fills the destination.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src),
               "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------- TMA (SASS: UTMALDG) + mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// Make freshly initialised barriers visible to the async proxy (the TMA unit).
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"  // sleep in hardware, do not spin
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// 3-D tiled TMA load: box (dim0 x dim1 x dim2) at coordinates (c0, c1, c2) of the tensor map into
// shared memory, completion counted in bytes on the mbarrier.
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* tmap, int c0, int c1, int c2,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
      : "memory");
}
// Order generic-proxy shared-memory writes before later async-proxy (TMA) writes to the same bytes.
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---------------------------------------------------------------- 16-byte chunk -> fp32 pairs
// A 16-byte chunk of the KV cache holds VEC elements; it is widened to VEC/2 packed fp32 pairs.
template <typename T>
struct Chunk;

template <>
struct Chunk<__nv_bfloat16> {
  static constexpr int VEC = 8;
  static constexpr int NP = 4;
  // bf16 -> fp32 is a 16-bit shift: low half (w << 16), high half (w & 0xffff0000).
  static __device__ __forceinline__ void widen(const uint4& v, u64 (&p)[4]) {
    p[0] = pack2(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xffff0000u));
    p[1] = pack2(__uint_as_float(v.y << 16), __uint_as_float(v.y & 0xffff0000u));
    p[2] = pack2(__uint_as_float(v.z << 16), __uint_as_float(v.z & 0xffff0000u));
    p[3] = pack2(__uint_as_float(v.w << 16), __uint_as_float(v.w & 0xffff0000u));
  }
  static __device__ __forceinline__ uint32_t narrow2(float lo, float hi) {
    __nv_bfloat162 r = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&r);
  }
};

template <>
struct Chunk<__half> {
  static constexpr int VEC = 8;
  static constexpr int NP = 4;
  static __device__ __forceinline__ u64 w2(uint32_t w) {
    float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    return pack2(f.x, f.y);
  }
  static __device__ __forceinline__ void widen(const uint4& v, u64 (&p)[4]) {
    p[0] = w2(v.x);
    p[1] = w2(v.y);
    p[2] = w2(v.z);
    p[3] = w2(v.w);
  }
  static __device__ __forceinline__ uint32_t narrow2(float lo, float hi) {
    __half2 r = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&r);
  }
};

template <>
struct Chunk<float> {
  static constexpr int VEC = 4;
  static constexpr int NP = 2;
  static __device__ __forceinline__ void widen(const uint4& v, u64 (&p)[2]) {
    p[0] = pack2(__uint_as_float(v.x), __uint_as_float(v.y));
    p[1] = pack2(__uint_as_float(v.z), __uint_as_float(v.w));
  }
};

// Store 4 consecutive fp32 results as 4 elements of T (16 B for fp32, 8 B for 16-bit types).
template <typename T>
__device__ __forceinline__ void store4(T* dst, float a, float b, float c, float d) {
  uint2 v;
  v.x = Chunk<T>::narrow2(a, b);
  v.y = Chunk<T>::narrow2(c, d);
  *reinterpret_cast<uint2*>(dst) = v;
}
template <>
__device__ __forceinline__ void store4<float>(float* dst, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(dst) = make_float4(a, b, c, d);
}

// Store 2 consecutive fp32 results as 2 elements of T.
template <typename T>
__device__ __forceinline__ void store2(T* dst, float a, float b) {
  *reinterpret_cast<uint32_t*>(dst) = Chunk<T>::narrow2(a, b);
}
template <>
__device__ __forceinline__ void store2<float>(float* dst, float a, float b) {
  *reinterpret_cast<float2*>(dst) = make_float2(a, b);
}

// Programmatic dependent launch: wait until the grid this one depends on has completed and its
// writes are visible (returns at once when the launch carried no programmatic dependency).
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

}  // namespace mvl
