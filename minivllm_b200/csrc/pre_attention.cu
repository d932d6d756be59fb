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
// cache rows.  HBM-bound byte work: 128-bit coalesced lanes, several (token, head) rows per warp,
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

// 16-byte lanes.  A row of one head is CH = D*es/16 chunks; lane l of a row's LR = CH/2 lanes owns
// chunk l of the first half and chunk l of the second half — exactly the element pairs the
// rotation mixes, so RoPE needs no cross-lane traffic and the RMS sum is a shuffle over LR lanes.
// A warp covers 32/LR rows at once (4 rows of 256 B for bf16, D = 128) with two 128-bit loads per
// lane in flight.
template <typename T, int D>
__global__ void __launch_bounds__(256) qk_norm_rope_store_kernel(const PreAttnParams p) {
  constexpr int VEC = Chunk<T>::VEC, NP = Chunk<T>::NP;
  constexpr int HALF = D / 2, LR = HALF / VEC, RW = 32 / LR;
  const int lane = threadIdx.x & 31, sub = lane / LR, l = lane % LR;
  const int rows_per_token = p.H + 2 * p.KVH;
  const long long total = p.num_tokens * rows_per_token;
  const long long row = ((long long)blockIdx.x * 8 + (threadIdx.x >> 5)) * RW + sub;
  const bool in_range = row < total;
  const long long rr = in_range ? row : total - 1;
  const long long t = rr / rows_per_token;
  const int r = (int)(rr - t * rows_per_token);
  const bool is_q = r < p.H, is_k = !is_q && r < p.H + p.KVH;
  const int head = is_q ? r : (is_k ? r - p.H : r - p.H - p.KVH);
  const T* src = (is_q ? reinterpret_cast<const T*>(p.q) + t * p.q_stride
                       : (is_k ? reinterpret_cast<const T*>(p.k) + t * p.k_stride
                               : reinterpret_cast<const T*>(p.v) + t * p.v_stride)) +
                 (long long)head * D + l * VEC;
  T* dst = nullptr;
  if (in_range) {
    if (is_q) {
      dst = reinterpret_cast<T*>(p.q_out) + (t * p.H + head) * D + l * VEC;
    } else if (p.slots != nullptr) {
      const long long s = p.slot_is_i64 ? reinterpret_cast<const long long*>(p.slots)[t]
                                        : (long long)reinterpret_cast<const int*>(p.slots)[t];
      if (s >= 0 && s < p.num_slots)  // -1: token has no slot (attention.py:29-30)
        dst = reinterpret_cast<T*>(is_k ? p.k_cache : p.v_cache) + (s * p.KVH + head) * D + l * VEC;
    }
  }
  const bool live = dst != nullptr;
  uint4 ra = make_uint4(0, 0, 0, 0), rb = ra;
  if (live) {
    ra = __ldg(reinterpret_cast<const uint4*>(src));
    rb = __ldg(reinterpret_cast<const uint4*>(src + HALF));
  }
  if (!is_q && !is_k) {  // v row: byte copy
    if (live) {
      *reinterpret_cast<uint4*>(dst) = ra;
      *reinterpret_cast<uint4*>(dst + HALF) = rb;
    }
    return;
  }
  float a[VEC], b[VEC];
  {
    u64 pa[NP], pb[NP];
    Chunk<T>::widen(ra, pa);
    Chunk<T>::widen(rb, pb);
#pragma unroll
    for (int e = 0; e < NP; ++e) {
      unpack2(pa[e], a[2 * e], a[2 * e + 1]);
      unpack2(pb[e], b[2 * e], b[2 * e + 1]);
    }
  }
  if (p.gamma != nullptr) {
    float ss = 0.f;
#pragma unroll
    for (int e = 0; e < VEC; ++e) ss += a[e] * a[e] + b[e] * b[e];
#pragma unroll
    for (int o = LR / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float inv = 1.0f / sqrtf(ss / (float)D + p.eps);
#pragma unroll
    for (int e = 0; e < VEC; e += 4) {
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(p.gamma + l * VEC + e));
      const float4 g2 = __ldg(reinterpret_cast<const float4*>(p.gamma + HALF + l * VEC + e));
      a[e] *= inv * g1.x; a[e + 1] *= inv * g1.y; a[e + 2] *= inv * g1.z; a[e + 3] *= inv * g1.w;
      b[e] *= inv * g2.x; b[e + 1] *= inv * g2.y; b[e + 2] *= inv * g2.z; b[e + 3] *= inv * g2.w;
    }
  }
  if (p.cos_sin != nullptr) {
    long long pos = p.pos_is_i64 ? reinterpret_cast<const long long*>(p.positions)[t]
                                 : (long long)reinterpret_cast<const int*>(p.positions)[t];
    pos = pos < 0 ? 0 : (pos >= p.max_position ? p.max_position - 1 : pos);
    const float* cs = p.cos_sin + pos * D + l * VEC;
#pragma unroll
    for (int e = 0; e < VEC; e += 4) {
      const float4 c = __ldg(reinterpret_cast<const float4*>(cs + e));
      const float4 s = __ldg(reinterpret_cast<const float4*>(cs + HALF + e));
      const float cc[4] = {c.x, c.y, c.z, c.w}, sn[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float x1 = a[e + u], x2 = b[e + u];
        a[e + u] = x1 * cc[u] - x2 * sn[u];
        b[e + u] = x1 * sn[u] + x2 * cc[u];
      }
    }
  }
  if (live) {
#pragma unroll
    for (int e = 0; e < VEC; e += 4) {
      store4<T>(dst + e, a[e], a[e + 1], a[e + 2], a[e + 3]);
      store4<T>(dst + HALF + e, b[e], b[e + 1], b[e + 2], b[e + 3]);
    }
  }
}

// Token-per-warp variant for large token counts (prefill).  The kernel above recomputes, for every
// 256-byte row, the row -> (token, head) division, the slot and position look-ups and reloads the
// 2 x 256 bytes of cos/sin that all rows of a token share; at 262,144 tokens that bookkeeping and
// the L1 traffic of the table (twice the payload) held it to 73 % of the HBM peak.  Here a warp
// owns a token: gamma lives in registers for the whole kernel, position / slot / cos / sin are
// fetched once per token, and the token's H + 2 KVH rows stream through in groups of U x (32/LR)
// rows with all 2U 128-bit loads of a lane issued before the first is used.
template <typename T, int D, int U>
__global__ void __launch_bounds__(256) qk_norm_rope_store_token_kernel(const PreAttnParams p) {
  constexpr int VEC = Chunk<T>::VEC, NP = Chunk<T>::NP;
  constexpr int HALF = D / 2, LR = HALF / VEC, RW = 32 / LR;
  const int lane = threadIdx.x & 31, sub = lane / LR, l = lane % LR;
  const int rows = p.H + 2 * p.KVH;
  const long long nwarps = (long long)gridDim.x * 8;
  const bool norm = p.gamma != nullptr, rope = p.cos_sin != nullptr;
  float g1[VEC], g2[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) {
    g1[e] = norm ? __ldg(p.gamma + l * VEC + e) : 1.f;
    g2[e] = norm ? __ldg(p.gamma + HALF + l * VEC + e) : 1.f;
  }
  for (long long t = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); t < p.num_tokens; t += nwarps) {
    float cc[VEC], sn[VEC];
    if (rope) {
      long long pos = p.pos_is_i64 ? reinterpret_cast<const long long*>(p.positions)[t]
                                   : (long long)reinterpret_cast<const int*>(p.positions)[t];
      pos = pos < 0 ? 0 : (pos >= p.max_position ? p.max_position - 1 : pos);
      const float* cs = p.cos_sin + pos * D + l * VEC;
#pragma unroll
      for (int e = 0; e < VEC; e += 4) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(cs + e));
        const float4 s4 = __ldg(reinterpret_cast<const float4*>(cs + HALF + e));
        cc[e] = c.x, cc[e + 1] = c.y, cc[e + 2] = c.z, cc[e + 3] = c.w;
        sn[e] = s4.x, sn[e + 1] = s4.y, sn[e + 2] = s4.z, sn[e + 3] = s4.w;
      }
    }
    long long slot = -1;
    if (p.slots != nullptr) {
      slot = p.slot_is_i64 ? reinterpret_cast<const long long*>(p.slots)[t]
                           : (long long)reinterpret_cast<const int*>(p.slots)[t];
      if (slot >= p.num_slots) slot = -1;  // -1: token has no slot (attention.py:29-30)
    }
    const T* qrow = reinterpret_cast<const T*>(p.q) + t * p.q_stride + l * VEC;
    const T* krow = reinterpret_cast<const T*>(p.k) + t * p.k_stride + l * VEC;
    const T* vrow = reinterpret_cast<const T*>(p.v) + t * p.v_stride + l * VEC;
    T* qdst = reinterpret_cast<T*>(p.q_out) + t * p.H * D + l * VEC;
    T* kdst = reinterpret_cast<T*>(p.k_cache) + slot * p.KVH * D + l * VEC;
    T* vdst = reinterpret_cast<T*>(p.v_cache) + slot * p.KVH * D + l * VEC;
    for (int r0 = 0; r0 < rows; r0 += RW * U) {
      uint4 ra[U], rb[U];
      T* dst[U];
      int kind[U];  // 0 q, 1 k, 2 v, -1 nothing to do
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int r = r0 + u * RW + sub;
        const T* src = qrow;
        kind[u] = -1, dst[u] = nullptr;
        if (r < p.H) {
          kind[u] = 0, src = qrow + r * D, dst[u] = qdst + r * D;
        } else if (r < rows && slot >= 0) {
          const bool is_k = r < p.H + p.KVH;
          const int head = is_k ? r - p.H : r - p.H - p.KVH;
          kind[u] = is_k ? 1 : 2, src = (is_k ? krow : vrow) + head * D, dst[u] = (is_k ? kdst : vdst) + head * D;
        }
        ra[u] = rb[u] = make_uint4(0, 0, 0, 0);
        if (kind[u] >= 0) {
          ra[u] = __ldg(reinterpret_cast<const uint4*>(src));
          rb[u] = __ldg(reinterpret_cast<const uint4*>(src + HALF));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        // every lane walks the same code (the RMS shuffles use the full mask and a group of rows can
        // mix k and v rows); v rows and idle lanes discard the arithmetic
        float a[VEC], b[VEC];
        {
          u64 pa[NP], pb[NP];
          Chunk<T>::widen(ra[u], pa);
          Chunk<T>::widen(rb[u], pb);
#pragma unroll
          for (int e = 0; e < NP; ++e) {
            unpack2(pa[e], a[2 * e], a[2 * e + 1]);
            unpack2(pb[e], b[2 * e], b[2 * e + 1]);
          }
        }
        if (norm) {  // the shuffles run for every lane of the warp (rows with nothing to do carry zeros)
          float ss = 0.f;
#pragma unroll
          for (int e = 0; e < VEC; ++e) ss += a[e] * a[e] + b[e] * b[e];
#pragma unroll
          for (int o = LR / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
          const float inv = 1.0f / sqrtf(ss / (float)D + p.eps);
#pragma unroll
          for (int e = 0; e < VEC; ++e) a[e] *= inv * g1[e], b[e] *= inv * g2[e];
        }
        if (rope) {
#pragma unroll
          for (int e = 0; e < VEC; ++e) {
            const float x1 = a[e], x2 = b[e];
            a[e] = x1 * cc[e] - x2 * sn[e];
            b[e] = x1 * sn[e] + x2 * cc[e];
          }
        }
        if (kind[u] == 2) {  // v row: byte copy
          *reinterpret_cast<uint4*>(dst[u]) = ra[u];
          *reinterpret_cast<uint4*>(dst[u] + HALF) = rb[u];
        } else if (kind[u] >= 0) {
#pragma unroll
          for (int e = 0; e < VEC; e += 4) {
            store4<T>(dst[u] + e, a[e], a[e + 1], a[e + 2], a[e + 3]);
            store4<T>(dst[u] + HALF + e, b[e], b[e + 1], b[e + 2], b[e + 3]);
          }
        }
      }
    }
  }
}

template <typename T>
static int launch_pre_t(const PreAttnParams& p, int head_dim, cudaStream_t s) {
  if (p.num_tokens >= 8192) {  // enough tokens for a warp per token on every SM (8 warps x 8 blocks x 148)
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      return (int)cudaGetLastError();
    int per_sm = 0;
    if (head_dim == 64)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, qk_norm_rope_store_token_kernel<T, 64, 4>, 256, 0);
    else if (head_dim == 128)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, qk_norm_rope_store_token_kernel<T, 128, 4>, 256, 0);
    else
      return -1000;
    const long long want = (p.num_tokens + 7) / 8, cap = (long long)sms * (per_sm > 0 ? per_sm : 1);
    const unsigned blocks = (unsigned)(want < cap ? want : cap);  // persistent: a warp strides over tokens
    if (head_dim == 64) qk_norm_rope_store_token_kernel<T, 64, 4><<<blocks, 256, 0, s>>>(p);
    else qk_norm_rope_store_token_kernel<T, 128, 4><<<blocks, 256, 0, s>>>(p);
    return (int)cudaGetLastError();
  }
  const long long rows = p.num_tokens * (p.H + 2 * p.KVH);
  const int rw = 32 / ((head_dim / 2) / Chunk<T>::VEC);  // rows per warp
  const unsigned blocks = (unsigned)((rows + 8 * rw - 1) / (8 * rw));
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
  if (bits & 15u) return -4;                                            // MVL_ERR_ALIGNMENT (16-byte lanes)
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
