// prefill_f32.cu — prefill attention for fp32 tensors (the reference's default dtype), CUDA cores.
//
// The tensor-core prefill kernels (prefill_tcgen05.cu, prefill.cu) take 16-bit inputs; an engine run
// in fp32 — what the reference does unless told otherwise — reaches
// Attention.forward's prefill branch with fp32 q/k/v, and north_star asks fp32 results within 1e-4 of
// the reference, which rules out TF32 products.  This kernel keeps everything in fp32:
//   CTA = 64 queries x 1 head, 128 threads as a 16 x 8 grid; thread (ty, tx) owns query rows
//   4ty..4ty+3, score columns tx + 8c (c < 8) of a 64-key tile and output columns 4tx + 32c.
//   Q, K, V tiles sit in shared memory (Q / K rows padded by 4 floats for conflict-free 128-bit
//   reads), filled with cp.async (zero-fill past the sequence / context end); S = Q.K^T and O += P.V
//   are register-tiled FFMA loops; the online softmax reduces each row over the 8 lanes that share
//   ty; P overlays the K tile.
// Dense (mvl_flash_prefill) or paged K/V with a cached prefix (mvl_flash_prefill_paged): same
// geometry as the tcgen05 kernels (query i of a sequence sits at position ctx - n + i).
// Compute-bound on the fp32 pipes (72 TFLOP/s peak); a correctness path, not a fast one.
#include "common.cuh"
#include "prefill.h"

namespace mvl {
namespace {

template <int D, bool PAGED>
__global__ void __launch_bounds__(128, 2) flash_prefill_f32_kernel(const PrefillParams p) {
  constexpr int BM = 64, BN = 64, LD = D + 4;   // Q / K rows padded by 4 floats (conflict-free 128-bit reads)
  constexpr int CPR = D / 4;                                   // 16-byte chunks per row
  extern __shared__ __align__(16) float smf[];
  float* sQ = smf;                 // BM x LD
  float* sK = sQ + BM * LD;        // BN x LD
  float* sV = sK + BN * LD;        // BN x D  (a warp reads one row at a time: no padding needed)
  float* sP = sK;                  // BM x BN probabilities overlay the K tile once S is done
  constexpr int PLD = LD;

  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
  const int mb = gridDim.x - 1 - blockIdx.x;  // heaviest (latest) query tiles first
  const int h = blockIdx.y, seq = blockIdx.z;
  const int kvh = h / (p.H / p.KVH);
  const int q_start = p.cu_seqlens[seq];
  const int n_new = p.cu_seqlens[seq + 1] - q_start;
  int ctx = n_new;
  if (PAGED) {
    ctx = p.ctx_is_i64 ? (int)reinterpret_cast<const long long*>(p.context_lens)[seq]
                       : reinterpret_cast<const int*>(p.context_lens)[seq];
    if (ctx < n_new) ctx = n_new;
  }
  const int prefix = ctx - n_new;
  const int m0 = mb * BM;
  if (m0 >= n_new) return;
  const int qa0 = prefix + m0;  // absolute position of the tile's first query
  const int ntiles = min((qa0 + BM - 1) / BN + 1, (ctx + BN - 1) / BN);
  const int* table = PAGED ? p.block_tables + (long long)seq * p.bt_stride : nullptr;

  // ---- Q tile (rows past the sequence end: zeros)
  for (int i = tid; i < BM * CPR; i += 128) {
    const int r = i / CPR, c = i % CPR;
    const bool valid = m0 + r < n_new;
    const float* src = reinterpret_cast<const float*>(p.q) + (long long)(q_start + (valid ? m0 + r : 0)) * p.q_stride +
                       h * D + c * 4;
    cp_async16(smem_u32(sQ + r * LD + c * 4), src, valid ? 16u : 0u);
  }
  auto load_kv = [&](int j) {
    for (int i = tid; i < BN * CPR; i += 128) {
      const int r = i / CPR, c = i % CPR;
      const int key = j * BN + r;
      bool valid = key < ctx;
      long long row = 0;
      if (PAGED) {
        if (valid) {
          const int e = __ldg(table + (key >> p.bs_shift));
          valid = e >= 0;
          row = ((long long)e << p.bs_shift) | (key & ((1 << p.bs_shift) - 1));
        }
      } else {
        row = q_start + key;
      }
      if (!valid) row = 0;
      const long long off = row * p.k_stride + kvh * D + c * 4;   // k_stride == v_stride in the paged case
      cp_async16(smem_u32(sK + r * LD + c * 4), reinterpret_cast<const float*>(p.k) + off, valid ? 16u : 0u);
      cp_async16(smem_u32(sV + r * D + c * 4),
                 reinterpret_cast<const float*>(p.v) + (PAGED ? off : row * p.v_stride + kvh * D + c * 4),
                 valid ? 16u : 0u);
    }
    cp_async_commit();
  };

  float o[4][D / 8];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < D / 8; ++c) o[r][c] = 0.f;
  float m_run[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY}, l_run[4] = {0.f, 0.f, 0.f, 0.f};

  for (int j = 0; j < ntiles; ++j) {
    __syncthreads();  // the previous tile's K, V and P are no longer read
    load_kv(j);
    cp_async_wait<0>();
    __syncthreads();

    // ---- S (4 rows x 8 columns per thread) = Q . K^T
    float s[4][8];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 8; ++c) s[r][c] = 0.f;
#pragma unroll 4
    for (int d = 0; d < D; d += 4) {
      float4 qv[4], kv[8];
#pragma unroll
      for (int r = 0; r < 4; ++r) qv[r] = *reinterpret_cast<const float4*>(sQ + (4 * ty + r) * LD + d);
#pragma unroll
      for (int c = 0; c < 8; ++c) kv[c] = *reinterpret_cast<const float4*>(sK + (tx + 8 * c) * LD + d);
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c)
          s[r][c] += qv[r].x * kv[c].x + qv[r].y * kv[c].y + qv[r].z * kv[c].z + qv[r].w * kv[c].w;
    }
    __syncthreads();  // every warp is done with the K tile: P may overwrite it
    // ---- scale (log2 units), causal mask, online softmax over the 8 lanes of a row
    const bool masked = (j + 1) * BN - 1 > qa0;  // the tile reaches past the first query's position
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int lim = qa0 + 4 * ty + r - j * BN;  // column c visible iff c <= lim
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float x = s[r][c] * p.qscale;
        if (masked && tx + 8 * c > lim) x = -INFINITY;
        s[r][c] = x;
        mx = fmaxf(mx, x);
      }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
      // key 0 of tile 0 is visible to every query, so the running maximum is finite from tile 0 on
      const float mn = fmaxf(m_run[r], mx);
      const float alpha = exp2f(m_run[r] - mn);  // first tile: exp2(-inf) = 0
      m_run[r] = mn;
      float rs = 0.f;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float pv = exp2f(s[r][c] - mn);
        rs += pv;
        sP[(4 * ty + r) * PLD + tx + 8 * c] = pv;
      }
      rs += __shfl_xor_sync(0xffffffffu, rs, 1);
      rs += __shfl_xor_sync(0xffffffffu, rs, 2);
      rs += __shfl_xor_sync(0xffffffffu, rs, 4);
      l_run[r] = l_run[r] * alpha + rs;
#pragma unroll
      for (int c = 0; c < D / 8; ++c) o[r][c] *= alpha;
    }
    __syncwarp();  // a row's P values are written and read by the 8 lanes of one warp
    // ---- O (4 rows x D/8 columns per thread) += P . V
#pragma unroll 2
    for (int k = 0; k < BN; ++k) {
      float pr[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) pr[r] = sP[(4 * ty + r) * PLD + k];
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        const float4 vv = *reinterpret_cast<const float4*>(sV + k * D + 4 * tx + 32 * c);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          o[r][4 * c + 0] += pr[r] * vv.x;
          o[r][4 * c + 1] += pr[r] * vv.y;
          o[r][4 * c + 2] += pr[r] * vv.z;
          o[r][4 * c + 3] += pr[r] * vv.w;
        }
      }
    }
  }
  cp_async_wait<0>();

#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int row = m0 + 4 * ty + r;
    if (row >= n_new) continue;
    const float inv = 1.f / l_run[r];
    float* out = reinterpret_cast<float*>(p.out) + ((long long)(q_start + row) * p.H + h) * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c)
      *reinterpret_cast<float4*>(out + 4 * tx + 32 * c) =
          make_float4(o[r][4 * c] * inv, o[r][4 * c + 1] * inv, o[r][4 * c + 2] * inv, o[r][4 * c + 3] * inv);
  }
}

template <int D, bool PAGED>
int launch_f32(const PrefillParams& p, int num_seqs, int max_seqlen, cudaStream_t s) {
  constexpr int SMEM = (2 * 64 * (D + 4) + 64 * D) * 4;   // Q, K (+P), V: two CTAs per SM
  auto kern = flash_prefill_f32_kernel<D, PAGED>;
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) configured[dev] = true;
  }
  dim3 grid((unsigned)((max_seqlen + 63) / 64), (unsigned)p.H, (unsigned)num_seqs);
  kern<<<grid, 128, SMEM, s>>>(p);
  return (int)cudaGetLastError();
}

}  // namespace

int launch_prefill_f32(const PrefillParams& p, int head_dim, int num_seqs, int max_seqlen, cudaStream_t s) {
  const bool paged = p.block_tables != nullptr;
  int r;
  if (head_dim == 64)
    r = paged ? launch_f32<64, true>(p, num_seqs, max_seqlen, s) : launch_f32<64, false>(p, num_seqs, max_seqlen, s);
  else
    r = paged ? launch_f32<128, true>(p, num_seqs, max_seqlen, s) : launch_f32<128, false>(p, num_seqs, max_seqlen, s);
  return r == 0 ? 0 : fail_cuda(r, "flash_prefill_f32 launch");
}

}  // namespace mvl
