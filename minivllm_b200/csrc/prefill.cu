// prefill.cu — varlen causal GQA flash attention for the prefill branch (sm_100a, 16-bit types).
//
// SURVEY.md §8(f) row N3.  Replaces the reference's Triton flash_attention_varlen_kernel and its
// wrapper flash_attention_prefill (reference src/myvllm/layers/attention.py:111-209, :212-280):
//   out[i, h] = softmax_{j <= i, same sequence}( scale * <q[i,h], k[j, h/G]> ) @ v[j, h/G]
// over the NEW tokens of each sequence, packed back to back and delimited by cu_seqlens
// (attention.py:136-138, causal mask :179-180, GQA map :133).  Like the reference it does not read
// cached-prefix pages.
//
// Mapping: one CTA = 64 queries of one sequence for one query head (grid = (query blocks, heads,
// sequences) as attention.py:267, heaviest query blocks first); 4 warps x 16 query rows.  K/V
// blocks of 64 keys stream through a double-buffered, XOR-swizzled cp.async stage and stop at the
// causal diagonal (the reference walks every block and masks).  Scores and P.V run on the tensor
// cores (mma.sync m16n8k16, fp32 accumulate) with the FlashAttention-2 register layout: q as A
// fragments, K via ldmatrix, online softmax on the accumulator rows (quad shuffles), P re-packed to
// 16-bit A fragments, V via ldmatrix.trans.  This mma.sync kernel serves batches of short sequences
// (max_seqlen <= 64) and non-positive scales; longer ones go to the tcgen05 kernels
// (prefill_tcgen05.cu, 3-4x faster), fp32 tensors to prefill_f32.cu.  This file also holds the two
// C entry points, mvl_flash_prefill and mvl_flash_prefill_paged, and the kernel choice.
#include "common.cuh"
#include "host.h"
#include "paged_decode_mma.cuh"
#include "prefill.h"

namespace mvl {

template <typename T, int D>
__global__ void __launch_bounds__(128, 2) flash_prefill_kernel(const PrefillParams p) {
  using M = MmaType<T>;
  constexpr int BM = 64, BN = 64, ROWB = D * 2, CPRW = ROWB / 16, KS = D / 16;
  constexpr int TILEB = BN * ROWB;  // one K (or V) block
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t sK = smem_u32(smem), sV = sK + 2 * TILEB;  // two stages each

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int mb = gridDim.x - 1 - blockIdx.x;  // heaviest (latest) query blocks first
  const int h = blockIdx.y, seq = blockIdx.z;
  const int kvh = h / (p.H / p.KVH);
  const int seq_start = p.cu_seqlens[seq];
  const int seq_len = p.cu_seqlens[seq + 1] - seq_start;
  const int m0 = mb * BM;
  if (m0 >= seq_len) return;

  // ---- q A fragments: rows m0 + warp*16 + {g, g+8}, 4 registers per k-step
  uint32_t qa[KS][4];
  {
    const int r0 = m0 + warp * 16 + g, r1 = r0 + 8;
    const T* q0 = reinterpret_cast<const T*>(p.q) + (long long)(seq_start + r0) * p.q_stride + h * D;
    const T* q1 = reinterpret_cast<const T*>(p.q) + (long long)(seq_start + r1) * p.q_stride + h * D;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int d = ks * 16 + 2 * t;
      qa[ks][0] = r0 < seq_len ? *reinterpret_cast<const uint32_t*>(q0 + d) : 0u;
      qa[ks][1] = r1 < seq_len ? *reinterpret_cast<const uint32_t*>(q1 + d) : 0u;
      qa[ks][2] = r0 < seq_len ? *reinterpret_cast<const uint32_t*>(q0 + d + 8) : 0u;
      qa[ks][3] = r1 < seq_len ? *reinterpret_cast<const uint32_t*>(q1 + d + 8) : 0u;
    }
  }

  // ---- K/V block loader: 128 threads x (BN*CPRW/128) 16-byte chunks, swizzled, zero-fill past the end
  const char* kg = reinterpret_cast<const char*>(p.k) + (long long)kvh * ROWB;
  const char* vg = reinterpret_cast<const char*>(p.v) + (long long)kvh * ROWB;
  auto load_block = [&](int nb, int stage) {
    constexpr int PER = BN * CPRW / 128;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int idx = threadIdx.x + i * 128;
      const int row = idx / CPRW, ch = idx % CPRW;
      const int key = nb * BN + row;
      const bool valid = key < seq_len;
      const long long tok = seq_start + (valid ? key : 0);
      const uint32_t off = row * ROWB + ((ch ^ (row & 7)) << 4);
      cp_async16(sK + stage * TILEB + off, kg + tok * p.k_stride * 2 + ch * 16, valid ? 16u : 0u);
      cp_async16(sV + stage * TILEB + off, vg + tok * p.v_stride * 2 + ch * 16, valid ? 16u : 0u);
    }
    cp_async_commit();
  };

  float o[D / 8][4];
#pragma unroll
  for (int j = 0; j < D / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m_r[2] = {-INFINITY, -INFINITY}, l_r[2] = {0.f, 0.f};  // rows g and g+8

  // causal: this CTA's queries see keys [0, min(seq_len, m0 + BM))
  const int kv_end = min(seq_len, m0 + BM);
  const int nblocks = (kv_end + BN - 1) / BN;
  const int mi = lane >> 3, mr = lane & 7;
  load_block(0, 0);
  for (int nb = 0; nb < nblocks; ++nb) {
    const int stage = nb & 1;
    if (nb + 1 < nblocks) load_block(nb + 1, stage ^ 1);
    else cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const uint32_t kt = sK + stage * TILEB, vt = sV + stage * TILEB;

    // ---- S (16 x 64) = Q . K^T
    float s[BN / 8][4];
#pragma unroll
    for (int j = 0; j < BN / 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int jp = 0; jp < BN / 16; ++jp) {
        // matrices: (keys 16jp + 0..7, chunk 2ks), (same, chunk 2ks+1), (keys +8..15, 2ks), (same, 2ks+1)
        const int key = jp * 16 + (mi >> 1) * 8 + mr, chk = 2 * ks + (mi & 1);
        uint32_t bfr[4];
        ldmatrix_x4(bfr, kt + key * ROWB + ((chk ^ (key & 7)) << 4));
        M::mma(s[2 * jp], qa[ks], bfr[0], bfr[1]);
        M::mma(s[2 * jp + 1], qa[ks], bfr[2], bfr[3]);
      }
    }
    // ---- scale, causal / length mask, online softmax on rows g (idx 0,1) and g+8 (idx 2,3)
    const int row0 = m0 + warp * 16 + g, row1 = row0 + 8;
    const bool diag = (nb + 1) * BN > m0;  // block touches the diagonal or the sequence end
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int j = 0; j < BN / 8; ++j) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float x = s[j][e] * p.qscale;
        if (diag) {
          const int key = nb * BN + j * 8 + 2 * t + (e & 1);
          const int row = (e < 2) ? row0 : row1;
          if (key > row || key >= seq_len) x = -INFINITY;
        }
        s[j][e] = x;
        mx[e >> 1] = fmaxf(mx[e >> 1], x);
      }
    }
    float alpha[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      // every row sees at least one unmasked key in every block it visits (key 0 in block 0, its
      // own position in the diagonal block), so the running maximum is finite from block 0 on
      const float mn = fmaxf(m_r[r], mx[r]);
      alpha[r] = fast_exp2(m_r[r] - mn);  // first block: exp2(-inf) = 0
      m_r[r] = mn;
      l_r[r] *= alpha[r];
      mx[r] = mn;
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int j = 0; j < BN / 8; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        s[j][e] = fast_exp2(s[j][e] - mx[e >> 1]);
        rs[e >> 1] += s[j][e];
      }
    l_r[0] += rs[0];
    l_r[1] += rs[1];
#pragma unroll
    for (int j = 0; j < D / 8; ++j) {
      o[j][0] *= alpha[0];
      o[j][1] *= alpha[0];
      o[j][2] *= alpha[1];
      o[j][3] *= alpha[1];
    }
    // ---- O (16 x D) += P (16 x 64) . V (64 x D)
#pragma unroll
    for (int kk = 0; kk < BN / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = M::pack(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = M::pack(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = M::pack(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = M::pack(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int jp = 0; jp < D / 16; ++jp) {
        // matrices (transposed on load): (keys 16kk+0..7, chunk 2jp), (keys +8..15, 2jp),
        //                                (keys 0..7, chunk 2jp+1), (keys +8..15, 2jp+1)
        const int key = kk * 16 + (mi & 1) * 8 + mr, chk = 2 * jp + (mi >> 1);
        uint32_t bfr[4];
        ldmatrix_x4_trans(bfr, vt + key * ROWB + ((chk ^ (key & 7)) << 4));
        M::mma(o[2 * jp], pa, bfr[0], bfr[1]);
        M::mma(o[2 * jp + 1], pa, bfr[2], bfr[3]);
      }
    }
    __syncthreads();  // everyone is done with this stage before it is refilled
  }
  cp_async_wait<0>();

  // ---- normalise and store rows g, g+8 (quad-reduced sums)
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_r[r] += __shfl_xor_sync(0xffffffffu, l_r[r], 1);
    l_r[r] += __shfl_xor_sync(0xffffffffu, l_r[r], 2);
  }
  const int row0 = m0 + warp * 16 + g, row1 = row0 + 8;
  T* out = reinterpret_cast<T*>(p.out);
#pragma unroll
  for (int j = 0; j < D / 8; ++j) {
    const int d = j * 8 + 2 * t;
    if (row0 < seq_len) {
      const float inv = 1.f / l_r[0];
      *reinterpret_cast<uint32_t*>(out + ((long long)(seq_start + row0) * p.H + h) * D + d) =
          M::pack(o[j][0] * inv, o[j][1] * inv);
    }
    if (row1 < seq_len) {
      const float inv = 1.f / l_r[1];
      *reinterpret_cast<uint32_t*>(out + ((long long)(seq_start + row1) * p.H + h) * D + d) =
          M::pack(o[j][2] * inv, o[j][3] * inv);
    }
  }
}

template <typename T, int D>
static int launch_prefill_td(const PrefillParams& p, int num_seqs, int max_seqlen, cudaStream_t s) {
  constexpr int SMEM = 4 * 64 * D * 2;  // two stages of K and of V
  auto kern = flash_prefill_kernel<T, D>;
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

}  // namespace mvl

extern "C" int mvl_flash_prefill(void* out, const void* q, const void* k, const void* v,
                                 int64_t q_token_stride, int64_t k_token_stride, int64_t v_token_stride,
                                 int64_t total_tokens, const int32_t* cu_seqlens, int num_seqs,
                                 int max_seqlen, int num_heads, int num_kv_heads, int head_dim, float scale,
                                 int elem_dtype, unsigned flags, void* stream) {
  if (elem_dtype < 0 || elem_dtype > 2) return -2;  // MVL_ERR_DTYPE
  const int es = elem_dtype == 0 ? 4 : 2;
  if (total_tokens < 0 || num_seqs < 0 || max_seqlen < 0 || num_heads <= 0 || num_kv_heads <= 0 || num_heads % num_kv_heads)
    return -3;
  if (head_dim != 64 && head_dim != 128) return -3;
  if (num_seqs == 0 || max_seqlen == 0 || total_tokens == 0) return 0;
  if (num_seqs > 65535 || num_heads > 65535) return -7;
  if (!out || !q || !k || !v || !cu_seqlens) return -1;
  if (q_token_stride < (int64_t)num_heads * head_dim || k_token_stride < (int64_t)num_kv_heads * head_dim ||
      v_token_stride < (int64_t)num_kv_heads * head_dim)
    return -3;
  const uintptr_t bits = (uintptr_t)out | (uintptr_t)q | (uintptr_t)k | (uintptr_t)v |
                         (uintptr_t)(q_token_stride * es) | (uintptr_t)(k_token_stride * es) |
                         (uintptr_t)(v_token_stride * es);
  if (bits & 15u) return -4;
  mvl::PrefillParams p;
  p.out = out; p.q = q; p.k = k; p.v = v; p.cu_seqlens = cu_seqlens;
  p.q_stride = q_token_stride; p.k_stride = k_token_stride; p.v_stride = v_token_stride;
  p.total_tokens = total_tokens;
  p.block_tables = nullptr; p.bt_stride = 0; p.context_lens = nullptr; p.ctx_is_i64 = 0;
  p.bs_shift = 0; p.piece_rows = 128; p.num_slots = 0;
  p.H = num_heads; p.KVH = num_kv_heads;
  p.qscale = scale * 1.4426950408889634f;
  cudaStream_t s = (cudaStream_t)stream;
  if (elem_dtype == 0) return mvl::launch_prefill_f32(p, head_dim, num_seqs, max_seqlen, s);
  // tcgen05 kernels: 128-query tiles, maximum taken on raw scores (needs scale > 0).  The pair kernel
  // (256 queries per CTA, ping-pong) wins from ~2k tokens per sequence, the one-tile kernel below
  // (1024: 582 vs 549 TFLOP/s; 4096: 904 vs 976); batches of short sequences keep mma.sync.
  const bool tc5_ok = scale > 0.f && total_tokens < (1ll << 31);
  if (!(flags & 1u) && tc5_ok && ((flags & 6u) || max_seqlen > 64)) {
    const bool single = (flags & 2u) || (!(flags & 4u) && max_seqlen < 2048);
    return mvl::launch_prefill_tcgen05(p, elem_dtype, head_dim, num_seqs, max_seqlen, single, s);
  }
  int e = -1000;
  if (elem_dtype == 2)
    e = head_dim == 64 ? mvl::launch_prefill_td<__nv_bfloat16, 64>(p, num_seqs, max_seqlen, s)
                       : mvl::launch_prefill_td<__nv_bfloat16, 128>(p, num_seqs, max_seqlen, s);
  else
    e = head_dim == 64 ? mvl::launch_prefill_td<__half, 64>(p, num_seqs, max_seqlen, s)
                       : mvl::launch_prefill_td<__half, 128>(p, num_seqs, max_seqlen, s);
  return e == 0 ? 0 : mvl::fail_cuda(e, "flash_prefill launch");
}

extern "C" int mvl_flash_prefill_paged(void* out, const void* q, int64_t q_token_stride, int64_t total_q_tokens,
                                       const int32_t* cu_seqlens_q, const void* k_cache, const void* v_cache,
                                       const int32_t* block_tables, int64_t block_table_stride,
                                       const void* context_lens, int ctx_is_i64, int num_seqs,
                                       int max_seqlen_q, int num_heads, int num_kv_heads, int head_dim,
                                       int block_size, int64_t num_blocks, float scale, int elem_dtype,
                                       unsigned flags, void* stream) {
  if (elem_dtype < 0 || elem_dtype > 2) return -2;  // MVL_ERR_DTYPE
  const int es = elem_dtype == 0 ? 4 : 2;
  if (total_q_tokens < 0 || num_seqs < 0 || max_seqlen_q < 0 || num_heads <= 0 || num_kv_heads <= 0 ||
      num_heads % num_kv_heads || num_blocks <= 0)
    return -3;
  if (head_dim != 64 && head_dim != 128) return -3;
  if (block_size < 16 || (block_size & (block_size - 1))) return -3;  // pages of 2^k >= 16 slots
  if (elem_dtype != 0 && !(scale > 0.f)) return -3;                    // tcgen05 kernels: raw-score maximum
  if (num_seqs == 0 || max_seqlen_q == 0 || total_q_tokens == 0) return 0;
  if (num_seqs > 65535 || num_heads > 65535) return -7;
  if (total_q_tokens >= (1ll << 31) || num_blocks * block_size >= (1ll << 31) - 256) return -7;
  if (!out || !q || !k_cache || !v_cache || !cu_seqlens_q || !block_tables || !context_lens) return -1;
  if (q_token_stride < (int64_t)num_heads * head_dim) return -3;
  const uintptr_t bits = (uintptr_t)out | (uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache |
                         (uintptr_t)(q_token_stride * es);
  if (bits & 15u) return -4;
  mvl::PrefillParams p;
  p.out = out; p.q = q; p.k = k_cache; p.v = v_cache; p.cu_seqlens = cu_seqlens_q;
  p.q_stride = q_token_stride;
  p.k_stride = p.v_stride = (long long)num_kv_heads * head_dim;
  p.total_tokens = total_q_tokens;
  p.block_tables = block_tables; p.bt_stride = block_table_stride;
  p.context_lens = context_lens; p.ctx_is_i64 = ctx_is_i64 ? 1 : 0;
  int shift = 0;
  while ((1 << shift) < block_size) ++shift;
  p.bs_shift = shift;
  p.piece_rows = block_size < 128 ? block_size : 128;
  p.num_slots = (int)(num_blocks * block_size);
  p.H = num_heads; p.KVH = num_kv_heads;
  p.qscale = scale * 1.4426950408889634f;
  if (elem_dtype == 0) return mvl::launch_prefill_f32(p, head_dim, num_seqs, max_seqlen_q, (cudaStream_t)stream);
  // with a cached prefix (or chunked prefill) a query tile walks many key tiles however few queries
  // there are, which is where the pair kernel's ping-pong pays: 8 x 1024 new tokens over 8k-token
  // contexts 1,087 vs 573 TFLOP/s.  One query tile per sequence leaves nothing to pair.
  const bool single = (flags & 2u) || (!(flags & 4u) && max_seqlen_q <= 128);
  return mvl::launch_prefill_tcgen05(p, elem_dtype, head_dim, num_seqs, max_seqlen_q, single,
                                     (cudaStream_t)stream);
}
