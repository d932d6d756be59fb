// prefill_tcgen05.cu — varlen causal GQA flash attention on the 5th-generation tensor cores.
//
// SURVEY.md §8(f) row N3, second version: the same contract as prefill.cu (reference
// src/myvllm/layers/attention.py:111-209 and :212-280) with both matrix products on tcgen05.mma:
//   S (128 x 128, fp32, TMEM)  = Q (128 x D, smem, K-major)  . K^T (128 keys x D, smem, K-major)
//   O (128 x D,  fp32, TMEM) += P (128 x 128, 16-bit, TMEM)  . V   (128 keys x D, smem, MN-major)
// One CTA = 128 queries of one sequence for one query head, 128 threads, thread t owns query row t
// (TMEM lane t): a whole score row arrives in its registers with tcgen05.ld, so the softmax needs
// no shuffles.  Q, K and V tiles are brought in by TMA (tensor maps over the packed tensors, 128-byte
// swizzle, rows past the tensor end zero-filled) and signal mbarriers; thread 0 issues the TMA loads
// and the MMAs, tcgen05.commit signals their completion.  P overwrites the first 64 columns of S
// (two 16-bit values per column) and is the A operand of the second product straight from TMEM.
// The running maximum is lazy: the reference exponent m_ref is refreshed (and O rescaled in TMEM)
// only when a row's maximum grows by more than 2^8, which after the first tile is rare.
// Two CTAs per SM (96 KB shared memory and 256 TMEM columns each): one CTA's softmax overlaps the
// other's MMAs.  K tile j+1 is loaded during softmax j, V tile j as soon as product j-1 has retired.
#include "common.cuh"
#include "paged_decode_mma.cuh"
#include "prefill.h"
#include "tcgen05.cuh"

namespace mvl {
namespace {

template <int D>
struct Tc5Cfg {
  static constexpr int BM = 128, BN = 128;
  static constexpr int PANELS = D / 64;             // 64 elements = one 128-byte swizzle row
  static constexpr int PANEL_BYTES = 128 * 128;     // 128 rows x 128 bytes
  static constexpr int TILE_BYTES = PANELS * PANEL_BYTES;
  static constexpr int OFF_K = TILE_BYTES, OFF_V = 2 * TILE_BYTES, OFF_BAR = 3 * TILE_BYTES;
  static constexpr int SMEM = OFF_BAR + 64 + 1024;  // + barriers / TMEM slot + alignment slack
  static constexpr int TMEM_COLS = 256;             // S: [0, 128), O: [128, 128 + D)
  static constexpr int O_COL = 128;
};

__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

template <typename T, int D>
__global__ void __launch_bounds__(128, 2)
flash_prefill_tcgen05_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                             const __grid_constant__ CUtensorMap mv, const PrefillParams p) {
  using C = Tc5Cfg<D>;
  using M = MmaType<T>;
  constexpr int FMT = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
  constexpr uint32_t IDESC_QK = tc5::instr_desc_f16(FMT, 0, 128, 128);
  constexpr uint32_t IDESC_PV = tc5::instr_desc_f16(FMT, 1, 128, D);
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // swizzle atoms are 1024-byte aligned
  const uint32_t sQ = base, sK = base + C::OFF_K, sV = base + C::OFF_V;
  const uint32_t bar_q = base + C::OFF_BAR, bar_k = bar_q + 8, bar_v = bar_q + 16, bar_s = bar_q + 24,
                 bar_o = bar_q + 32, tmem_slot = bar_q + 40;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int mb = gridDim.x - 1 - blockIdx.x;  // heaviest (latest) query tiles first
  const int h = blockIdx.y, seq = blockIdx.z;
  const int kvh = h / (p.H / p.KVH);
  const int seq_start = p.cu_seqlens[seq];
  const int seq_len = p.cu_seqlens[seq + 1] - seq_start;
  const int m0 = mb * C::BM;
  if (m0 >= seq_len) return;
  const int ntiles = mb + 1;  // key tiles 0..mb; the last one holds the diagonal

  if (warp == 0) {
    tc5::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc5::tmem_relinquish();
  }
  if (tid == 32) {
    mbar_init(bar_q, 1), mbar_init(bar_k, 1), mbar_init(bar_v, 1), mbar_init(bar_s, 1), mbar_init(bar_o, 1);
    mbar_fence_init();
  }
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = lds_u32(tmem_slot);
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);

  auto load_tile = [&](const CUtensorMap* map, uint32_t dst, int head, int token, uint32_t bar) {
    mbar_expect_tx(bar, C::TILE_BYTES);
#pragma unroll
    for (int pn = 0; pn < C::PANELS; ++pn) tma_load_3d(dst + pn * C::PANEL_BYTES, map, pn * 64, head, token, bar);
  };
  auto issue_qk = [&]() {  // S = Q . K^T: D/16 steps of K = 16 (32 bytes along the swizzled row)
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      const uint32_t off = (ks >> 2) * C::PANEL_BYTES + (ks & 3) * 32;
      tc5::mma_ss(tmem, tc5::smem_desc_sw128(sQ + off, 16, 1024), tc5::smem_desc_sw128(sK + off, 16, 1024),
                  IDESC_QK, ks > 0);
    }
    tc5::commit(bar_s);
  };
  auto issue_pv = [&](int j) {  // O += P . V: 8 steps of 16 keys (8 TMEM columns of P, 16 rows of V)
#pragma unroll
    for (int kk = 0; kk < C::BN / 16; ++kk)
      tc5::mma_ts(tmem + C::O_COL, tmem + kk * 8, tc5::smem_desc_sw128(sV + kk * 2048, C::PANEL_BYTES, 1024),
                  IDESC_PV, (j > 0 || kk > 0) ? 1u : 0u);
    tc5::commit(bar_o);
  };

  if (tid == 0) {
    load_tile(&mq, sQ, h, seq_start + m0, bar_q);
    load_tile(&mk, sK, kvh, seq_start, bar_k);
    load_tile(&mv, sV, kvh, seq_start, bar_v);
    tc5::mbar_wait_bounded(bar_q, 0);
    tc5::mbar_wait_bounded(bar_k, 0);
    tc5::fence_after_sync();
    issue_qk();
  }

  float m_ref = 0.f, l = 0.f;
  const float qs = p.qscale;
  for (int j = 0; j < ntiles; ++j) {
    tc5::mbar_wait_bounded(bar_s, j & 1);  // S_j is complete, and so is every MMA issued before it
    tc5::fence_after_sync();
    if (tid == 0) {
      if (j > 0) load_tile(&mv, sV, kvh, seq_start + j * C::BN, bar_v);               // product j-1 has retired
      if (j + 1 < ntiles) load_tile(&mk, sK, kvh, seq_start + (j + 1) * C::BN, bar_k);  // S_j no longer reads K_j
    }
    uint32_t su[128];
#pragma unroll
    for (int c = 0; c < 4; ++c) tc5::tmem_ld32(lane_base + c * 32, su + c * 32);
    tc5::tmem_wait_ld();
    float mx = -INFINITY;
    if (j == ntiles - 1) {  // diagonal tile: key (kv0 + c) is visible to query (m0 + tid) iff c <= tid
#pragma unroll
      for (int c = 0; c < 128; ++c) {
        if (c > tid) su[c] = 0xff800000u;
        mx = fmaxf(mx, __uint_as_float(su[c]));
      }
    } else {
#pragma unroll
      for (int c = 0; c < 128; ++c) mx = fmaxf(mx, __uint_as_float(su[c]));
    }
    bool grow = false;
    if (j == 0) m_ref = mx;
    else grow = (mx - m_ref) * qs > 8.f;
    if (__any_sync(0xffffffffu, grow)) {  // warp-uniform: the TMEM moves are warp-collective
      float alpha = 1.f;
      if (grow) {
        alpha = fast_exp2((m_ref - mx) * qs);
        m_ref = mx;
        l *= alpha;
      }
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        tc5::tmem_ld32(lane_base + C::O_COL + c * 32, o);
        tc5::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
        tc5::tmem_st32(lane_base + C::O_COL + c * 32, o);
      }
    }
    const float neg = -m_ref * qs;
    float sum = 0.f;
    uint32_t pk[64];
#pragma unroll
    for (int c = 0; c < 64; ++c) {
      const float p0 = fast_exp2(fmaf(__uint_as_float(su[2 * c]), qs, neg));
      const float p1 = fast_exp2(fmaf(__uint_as_float(su[2 * c + 1]), qs, neg));
      sum += p0 + p1;
      pk[c] = M::pack(p0, p1);
    }
    l += sum;
    tc5::tmem_st32(lane_base, pk);
    tc5::tmem_st32(lane_base + 32, pk + 32);
    tc5::tmem_wait_st();
    tc5::fence_before_sync();
    __syncthreads();  // every row of P_j (and any rescaled O) is in TMEM
    if (tid == 0) {
      tc5::fence_after_sync();
      tc5::mbar_wait_bounded(bar_v, j & 1);
      issue_pv(j);
      if (j + 1 < ntiles) {
        tc5::mbar_wait_bounded(bar_k, (j + 1) & 1);
        tc5::fence_after_sync();
        issue_qk();  // queued behind product j, which still reads P_j from the columns S_{j+1} overwrites
      }
    }
  }
  tc5::mbar_wait_bounded(bar_o, (ntiles - 1) & 1);
  tc5::fence_after_sync();

  // ---- O / l -> out: thread t writes row m0 + t
  const int row = m0 + tid;
  const float inv = 1.f / l;
  T* out = reinterpret_cast<T*>(p.out) + ((long long)(seq_start + row) * p.H + h) * D;
#pragma unroll
  for (int c = 0; c < D / 32; ++c) {
    uint32_t o[32];
    tc5::tmem_ld32(lane_base + C::O_COL + c * 32, o);
    tc5::tmem_wait_ld();
    if (row < seq_len) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 w;
        w.x = M::pack(__uint_as_float(o[8 * i + 0]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
        w.y = M::pack(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
        w.z = M::pack(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
        w.w = M::pack(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
        *reinterpret_cast<uint4*>(out + c * 32 + i * 8) = w;
      }
    }
  }
  tc5::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc5::tmem_dealloc(tmem, C::TMEM_COLS);
}

template <typename T, int D>
int launch_td(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const PrefillParams& p,
              int num_seqs, int max_seqlen, cudaStream_t s) {
  using C = Tc5Cfg<D>;
  auto kern = flash_prefill_tcgen05_kernel<T, D>;
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) configured[dev] = true;
  }
  dim3 grid((unsigned)((max_seqlen + C::BM - 1) / C::BM), (unsigned)p.H, (unsigned)num_seqs);
  kern<<<grid, 128, C::SMEM, s>>>(mq, mk, mv, p);
  return (int)cudaGetLastError();
}

}  // namespace

int launch_prefill_tcgen05(const PrefillParams& p, int dtype, int head_dim, int num_seqs, int max_seqlen,
                           cudaStream_t s) {
  CUtensorMap mq, mk, mv;
  int r;
  if ((r = make_packed_map(&mq, p.q, dtype, head_dim, p.H, p.total_tokens, p.q_stride, 128)) != 0)
    return fail_text("prefill tensor map (q)", r);
  if ((r = make_packed_map(&mk, p.k, dtype, head_dim, p.KVH, p.total_tokens, p.k_stride, 128)) != 0)
    return fail_text("prefill tensor map (k)", r);
  if ((r = make_packed_map(&mv, p.v, dtype, head_dim, p.KVH, p.total_tokens, p.v_stride, 128)) != 0)
    return fail_text("prefill tensor map (v)", r);
  if (dtype == 2)
    r = head_dim == 64 ? launch_td<__nv_bfloat16, 64>(mq, mk, mv, p, num_seqs, max_seqlen, s)
                       : launch_td<__nv_bfloat16, 128>(mq, mk, mv, p, num_seqs, max_seqlen, s);
  else
    r = head_dim == 64 ? launch_td<__half, 64>(mq, mk, mv, p, num_seqs, max_seqlen, s)
                       : launch_td<__half, 128>(mq, mk, mv, p, num_seqs, max_seqlen, s);
  return r == 0 ? 0 : fail_cuda(r, "flash_prefill_tcgen05 launch");
}

}  // namespace mvl
