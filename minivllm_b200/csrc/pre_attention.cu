// pre_attention.cu — q/k RMSNorm + rotary embedding + KV-cache store in ONE pass (sm_100a).
//
// the neighbours of the hot path.  Between the QKV projection and
// attention the reference runs, per layer (reference src/myvllm/models/qwen3.py:104-110):
//   q = rms_norm(q); k = rms_norm(k)          layers/layernorm.py:13-15, x / sqrt(mean(x^2)+eps) * gamma
//   q, k = rotary_emb(positions, q, k)        layers/rotary_embedding.py:74-83 -> :4-23 (half split:
//                                             out1 = x1*cos - x2*sin, out2 = x1*sin + x2*cos)
//   store_kvcache(k, v, ...)                  layers/attention.py:498-510
// i.e. two torch.compile'd norm kernels, a gather + rotate, two contiguous() copies and the Triton
// store.  Here one kernel reads each q/k/v row once and writes q (normalised, rotated) and the
// cache rows.  HBM-bound byte work: one warp per (token, head) row, coalesced 4- or 8-byte lanes,
// fp32 arithmetic, one rounding on store.
#include "common.cuh"
#include "host.h"

namespace mvl {

struct PreAttnParams {
  void* q_out;
  const void* q;
  const void* k;
  const void* v;
  long long q_stride, k_stride, v_stride;  // token strides, elements
  const float* gamma;                      // (D) or nullptr: no RMSNorm (qkv_bias models)
  float eps;
  const float* cos_sin;                    // (max_position, D): [cos(D/2) | sin(D/2)], fp32
  long long max_position;
  const void* positions;
  int pos_is_i64;
  void* k_cache;
  void* v_cache;
  const void* slots;                       // nullptr: no cache write
  int slot_is_i64;
  long long num_slots, num_tokens;
  int H, KVH;
};

template <typename T> struct Pair;  // two adjacent elements <-> float2
template <> struct Pair<float> {
  static __device__ __forceinline__ float2 ld(const float* p) { return *reinterpret_cast<const float2*>(p); }
  static __device__ __forceinline__ void st(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }
};
template <> struct Pair<__nv_bfloat16> {
  static __device__ __forceinline__ float2 ld(const __nv_bfloat16* p) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float2 v) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(v.x, v.y);
  }
};
template <> struct Pair<__half> {
  static __device__ __forceinline__ float2 ld(const __half* p) {
    return __half22float2(*reinterpret_cast<const __half2*>(p));
  }
  static __device__ __forceinline__ void st(__half* p, float2 v) {
    *reinterpret_cast<__half2*>(p) = __floats2half2_rn(v.x, v.y);
  }
};

// One warp per row.  Lane l (< D/4) owns elements {2l, 2l+1} of the first half and the same two of
// the second half — exactly the pairs the rotation mixes, so RoPE needs no cross-lane traffic.
template <typename T, int D>
__global__ void __launch_bounds__(256) qk_norm_rope_store_kernel(const PreAttnParams p) {
  constexpr int HALF = D / 2;
  const int lane = threadIdx.x & 31;
  const int rows_per_token = p.H + 2 * p.KVH;
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= p.num_tokens * rows_per_token) return;
  const long long t = row / rows_per_token;
  const int r = (int)(row - t * rows_per_token);
  const bool active = lane < HALF / 2;
  const int i0 = 2 * lane;  // index inside a half

  const bool is_q = r < p.H, is_k = !is_q && r < p.H + p.KVH;
  const int head = is_q ? r : (is_k ? r - p.H : r - p.H - p.KVH);
  const T* src = is_q ? reinterpret_cast<const T*>(p.q) + t * p.q_stride
                      : (is_k ? reinterpret_cast<const T*>(p.k) + t * p.k_stride
                              : reinterpret_cast<const T*>(p.v) + t * p.v_stride);
  src += (long long)head * D;

  T* dst = nullptr;
  if (is_q) {
    dst = reinterpret_cast<T*>(p.q_out) + (t * p.H + head) * D;
  } else if (p.slots != nullptr) {
    const long long s = p.slot_is_i64 ? reinterpret_cast<const long long*>(p.slots)[t]
                                      : (long long)reinterpret_cast<const int*>(p.slots)[t];
    if (s >= 0 && s < p.num_slots)  // -1: token has no slot (attention.py:29-30)
      dst = reinterpret_cast<T*>(is_k ? p.k_cache : p.v_cache) + (s * p.KVH + head) * D;
  }
  if (dst == nullptr) return;  // warp-uniform

  float2 a = make_float2(0.f, 0.f), b = make_float2(0.f, 0.f);  // first-half pair, second-half pair
  if (active) {
    a = Pair<T>::ld(src + i0);
    b = Pair<T>::ld(src + HALF + i0);
  }
  if (is_q || is_k) {
    if (p.gamma != nullptr) {
      float ss = a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
      const float rms = sqrtf(ss / (float)D + p.eps);
      if (active) {
        const float2 g1 = *reinterpret_cast<const float2*>(p.gamma + i0);
        const float2 g2 = *reinterpret_cast<const float2*>(p.gamma + HALF + i0);
        a.x = a.x / rms * g1.x;
        a.y = a.y / rms * g1.y;
        b.x = b.x / rms * g2.x;
        b.y = b.y / rms * g2.y;
      }
    }
    if (p.cos_sin != nullptr && active) {
      long long pos = p.pos_is_i64 ? reinterpret_cast<const long long*>(p.positions)[t]
                                   : (long long)reinterpret_cast<const int*>(p.positions)[t];
      pos = pos < 0 ? 0 : (pos >= p.max_position ? p.max_position - 1 : pos);
      const float2 c = *reinterpret_cast<const float2*>(p.cos_sin + pos * D + i0);
      const float2 s = *reinterpret_cast<const float2*>(p.cos_sin + pos * D + HALF + i0);
      const float2 x1 = a, x2 = b;
      a.x = x1.x * c.x - x2.x * s.x;
      a.y = x1.y * c.y - x2.y * s.y;
      b.x = x1.x * s.x + x2.x * c.x;
      b.y = x1.y * s.y + x2.y * c.y;
    }
  }
  if (active) {
    Pair<T>::st(dst + i0, a);
    Pair<T>::st(dst + HALF + i0, b);
  }
}

template <typename T>
static int launch_pre_t(const PreAttnParams& p, int head_dim, cudaStream_t s) {
  const long long rows = p.num_tokens * (p.H + 2 * p.KVH);
  const unsigned blocks = (unsigned)((rows + 7) / 8);
  switch (head_dim) {
    case 64: qk_norm_rope_store_kernel<T, 64><<<blocks, 256, 0, s>>>(p); break;
    case 128: qk_norm_rope_store_kernel<T, 128><<<blocks, 256, 0, s>>>(p); break;
    default: return -1000;
  }
  return (int)cudaGetLastError();
}

int launch_pre_attention(const PreAttnParams& p, int head_dim, int dtype, cudaStream_t s) {
  switch (dtype) {
    case 0: return launch_pre_t<float>(p, head_dim, s);
    case 1: return launch_pre_t<__half>(p, head_dim, s);
    case 2: return launch_pre_t<__nv_bfloat16>(p, head_dim, s);
  }
  return -1000;
}

}  // namespace mvl

extern "C" int mvl_qk_norm_rope_store(void* q_out, const void* q, const void* k, const void* v,
                                      int64_t q_token_stride, int64_t k_token_stride,
                                      int64_t v_token_stride, const float* gamma, float eps,
                                      const float* cos_sin_cache, int64_t max_position,
                                      const void* positions, int pos_is_i64, void* k_cache,
                                      void* v_cache, const void* slot_mapping, int slot_is_i64,
                                      int64_t num_slots, int64_t num_tokens, int num_heads,
                                      int num_kv_heads, int head_dim, int elem_dtype, void* stream) {
  if (elem_dtype < 0 || elem_dtype > 2) return -2;                      // MVL_ERR_DTYPE
  if (num_tokens < 0 || num_heads <= 0 || num_kv_heads <= 0) return -3;  // MVL_ERR_SHAPE
  if (head_dim != 64 && head_dim != 128) return -3;
  if (num_tokens == 0) return 0;
  if (!q_out || !q || !k || !v) return -1;                              // MVL_ERR_NULL_POINTER
  if (cos_sin_cache && (!positions || max_position <= 0)) return -1;
  if (slot_mapping && (!k_cache || !v_cache)) return -1;
  const int es = elem_dtype == 0 ? 4 : 2;
  if (q_token_stride < (int64_t)num_heads * head_dim || k_token_stride < (int64_t)num_kv_heads * head_dim ||
      v_token_stride < (int64_t)num_kv_heads * head_dim)
    return -3;
  const uintptr_t bits = (uintptr_t)q_out | (uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)k_cache |
                         (uintptr_t)v_cache | (uintptr_t)gamma | (uintptr_t)cos_sin_cache |
                         (uintptr_t)(q_token_stride * es) | (uintptr_t)(k_token_stride * es) |
                         (uintptr_t)(v_token_stride * es);
  if (bits & 7u) return -4;                                             // MVL_ERR_ALIGNMENT (8-byte lanes)
  mvl::PreAttnParams p;
  p.q_out = q_out; p.q = q; p.k = k; p.v = v;
  p.q_stride = q_token_stride; p.k_stride = k_token_stride; p.v_stride = v_token_stride;
  p.gamma = gamma; p.eps = eps; p.cos_sin = cos_sin_cache; p.max_position = max_position;
  p.positions = positions; p.pos_is_i64 = pos_is_i64 ? 1 : 0;
  p.k_cache = k_cache; p.v_cache = v_cache; p.slots = slot_mapping; p.slot_is_i64 = slot_is_i64 ? 1 : 0;
  p.num_slots = num_slots; p.num_tokens = num_tokens; p.H = num_heads; p.KVH = num_kv_heads;
  const int e = mvl::launch_pre_attention(p, head_dim, elem_dtype, (cudaStream_t)stream);
  if (e == -1000) return -3;
  return e == 0 ? 0 : -6;                                               // MVL_ERR_CUDA
}
