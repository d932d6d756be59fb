// tcgen05.cuh — PTX wrappers for the 5th-generation tensor cores of sm_100a: tensor-memory (TMEM)
// allocation, tcgen05.mma with shared-memory descriptors (or the A operand in TMEM), commit to an
// mbarrier, and the 32x32b TMEM <-> register moves.  Encodings of the two descriptors follow the
// PTX ISA tables for tcgen05 (matrix descriptor, instruction descriptor, kind::f16).
#pragma once
#include <cstdint>

namespace mvl {
namespace tc5 {

// ---------------------------------------------------------------- TMEM allocation (one warp, .sync.aligned)
__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// The mbarrier receives one arrival when every tcgen05.mma this thread issued before has completed.
__device__ __forceinline__ void commit(uint32_t bar_smem) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_smem)
               : "memory");
}

// ---------------------------------------------------------------- descriptors
// Shared-memory matrix descriptor, 128-byte swizzle: start address, leading / stride byte offsets
// (all >> 4), descriptor version 1 (bit 46), swizzle mode 2 = 128B (bits 61..63).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) |
         ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// The same descriptor split in two 32-bit words, so that the issuing thread's loop only adds a
// byte offset >> 4 to the low word (one uniform add per MMA: the issue loop must stay well under the
// 64 clocks a 128 x 128 x 16 MMA takes, or the tensor pipe waits for its next instruction).
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t addr, uint32_t lbo_bytes) {
  return ((addr & 0x3FFFFu) >> 4) | ((lbo_bytes >> 4) << 16);
}
__host__ __device__ constexpr uint32_t smem_desc_hi_sw128(uint32_t sbo_bytes) {
  return (sbo_bytes >> 4) | (1u << 14) | (2u << 29);
}
__device__ __forceinline__ uint64_t desc64(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }

// Instruction descriptor of kind::f16: fp32 accumulate (bits 4..5 = 1), A/B format (0 = f16, 1 = bf16)
// at bits 7..9 / 10..12, A/B major (0 = K, 1 = MN) at bits 15 / 16, N >> 3 at 17..22, M >> 4 at 24..28.
__host__ __device__ constexpr uint32_t instr_desc_f16(int fmt, int b_mn_major, int m, int n) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// D[tmem] (+)= A[smem] . B[smem]
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                       uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]   (A: row m in lane m, 16-bit elements packed two per column)
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                       uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ---------------------------------------------------------------- TMEM <-> registers, 32 lanes x 32 columns
#define MVL_O4(r, i) "=r"(r[i]), "=r"(r[i + 1]), "=r"(r[i + 2]), "=r"(r[i + 3])
#define MVL_I4(r, i) "r"(r[i]), "r"(r[i + 1]), "r"(r[i + 2]), "r"(r[i + 3])
// Thread t of warp w reads lane 32*(w%4)+t, columns [col, col+32) of taddr (lane field = 32*(w%4)).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : MVL_O4(r, 0), MVL_O4(r, 4), MVL_O4(r, 8), MVL_O4(r, 12), MVL_O4(r, 16), MVL_O4(r, 20), MVL_O4(r, 24),
        MVL_O4(r, 28)
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};" ::MVL_I4(r, 0),
      MVL_I4(r, 4), MVL_I4(r, 8), MVL_I4(r, 12), MVL_I4(r, 16), MVL_I4(r, 20), MVL_I4(r, 24), MVL_I4(r, 28),
      "r"(taddr)
      : "memory");
}
#undef MVL_O4
#undef MVL_I4
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Start fetching a tensor map (a __grid_constant__ kernel parameter) before its first TMA use.
__device__ __forceinline__ void prefetch_tensormap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Non-blocking: has the phase with this parity completed?
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

// mbarrier wait that traps instead of hanging the GPU if the barrier never completes (a lost TMA
// or MMA completion is a programming error; ~2^28 polls is seconds, far beyond any real wait).
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
  // the loop lives inside one asm block: a C-level spin loop here keeps ptxas from giving the
  // softmax warpgroups the registers they obtain with setmaxnreg (measured: it then spills)
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .u32 spins;\n"
      "mov.u32 spins, 0;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"   // suspend-time hint: the
      "@p bra DONE_%=;\n"                                                   // thread sleeps in hardware
      "add.u32 spins, spins, 1;\n"                                         // instead of polling
      "setp.gt.u32 p, spins, 0x100000;\n"
      "@p trap;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

}  // namespace tc5
}  // namespace mvl
