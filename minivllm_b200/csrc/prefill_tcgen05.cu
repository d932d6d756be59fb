// prefill_tcgen05.cu — varlen causal GQA flash attention on the 5th-generation tensor cores.
//
// the contract of reference src/myvllm/layers/attention.py:111-209 and
// :212-280 (and, with PAGED, of the cached-prefix case the reference sets up in
// model_runner.py:209-215 but does not compute, A-8) with both matrix products on tcgen05.mma:
//   S (128 x 128, fp32, TMEM)  = Q (128 x D, smem, K-major)  . K^T (128 keys x D, smem, K-major)
//   O (128 x D,  fp32, TMEM) += P (128 x 128, 16-bit, TMEM)  . V   (128 keys x D, smem, MN-major)
// Thread t of a 128-thread group owns query row t (TMEM lane t): a whole score row arrives in its
// registers with tcgen05.ld, so the softmax needs no shuffles.  Q, K and V tiles are brought in by TMA
// (tensor maps over the packed tensors or over the paged cache, 128-byte swizzle, rows outside the
// map zero-filled) and signal mbarriers; tcgen05.commit signals MMA completion.  P overwrites the first
// 64 columns of S (two 16-bit values per column) and is the A operand of the second product straight
// from TMEM.  The running maximum is lazy: the reference exponent m_ref is refreshed (and O rescaled
// in TMEM) only when a row's maximum grows by more than 2^8, which after the first tile is rare.
// Two kernels share the softmax step and the tile loads:
//   flash_prefill_tcgen05_kernel       one 128-query tile per CTA, 128 threads, two CTAs per SM (96 KB
//                                      shared memory, 256 TMEM columns each): one CTA's softmax overlaps
//                                      the other's MMAs; warp 0 issues loads, thread 0 the MMAs
//   flash_prefill_tcgen05_pair_kernel  two tiles per CTA in ping-pong, one CTA per SM, dedicated TMA and
//                                      MMA warps, setmaxnreg (described at its definition)
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

__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// One softmax step of one query row (thread = row = TMEM lane): S row -> registers, causal mask (key
// column c of this tile is visible iff c <= lim; lim >= 127: nothing to mask), lazy maximum (O rescaled in TMEM only when the row maximum
// grew by more than 2^8), P = exp2((S - m_ref) * qs) packed to 16 bits over the first 64 S columns.
// Ends with tcgen05.wait::st: the caller fences and signals.
template <typename T, int D>
__device__ __forceinline__ void softmax_step(uint32_t s_addr, uint32_t o_addr, bool first, int lim, float qs,
                                             float& m_ref, float& l, uint32_t half_bar = 0) {
  using M = MmaType<T>;
  uint32_t su[128];
#pragma unroll
  for (int c = 0; c < 4; ++c) tc5::tmem_ld32(s_addr + c * 32, su + c * 32);
  tc5::tmem_wait_ld();
  if (lim < 127) {
#pragma unroll
    for (int c = 0; c < 128; ++c)
      if (c > lim) su[c] = 0xff800000u;
  }
  float mxa[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // four chains: the row maximum is latency-bound
#pragma unroll
  for (int c = 0; c < 128; c += 8)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      mxa[k] = fmax3(mxa[k], __uint_as_float(su[c + 2 * k]), __uint_as_float(su[c + 2 * k + 1]));
  const float mx = fmaxf(fmaxf(mxa[0], mxa[1]), fmaxf(mxa[2], mxa[3]));
  bool grow = false;
  if (first) m_ref = mx;
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
      tc5::tmem_ld32(o_addr + c * 32, o);
      tc5::tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
      tc5::tmem_st32(o_addr + c * 32, o);
    }
  }
  const float neg = -m_ref * qs;
  const u64 qs2 = pack2(qs, qs), neg2 = pack2(neg, neg);
  u64 acc[4] = {0ull, 0ull, 0ull, 0ull};
  // P in two halves of 64 keys: with half_bar set, the first half is published as soon as it is in
  // TMEM, so that the first four MMAs of P.V run while the second half's exponentials do
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    uint32_t pk[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      const int cc = half * 32 + c;
      float x0, x1;
      unpack2(ffma2(pack2(__uint_as_float(su[2 * cc]), __uint_as_float(su[2 * cc + 1])), qs2, neg2), x0, x1);
      // MUFU.EX2 for every element: a polynomial 2^x on the FMA pipe for 1, 2 or 3 of every 4 column
      // pairs was measured slower (1175 / 1104 / 1023 vs 1202 TFLOP/s at 2 x 16k tokens)
      const float p0 = fast_exp2(x0), p1 = fast_exp2(x1);
      acc[c & 3] = fadd2(acc[c & 3], pack2(p0, p1));
      pk[c] = M::pack(p0, p1);
    }
    tc5::tmem_st32(s_addr + half * 32, pk);
    if (half == 0 && half_bar != 0) {
      tc5::tmem_wait_st();
      tc5::fence_before_sync();
      tc5::mbar_arrive(half_bar);
    }
  }
  l += hsum2(fadd2(fadd2(acc[0], acc[1]), fadd2(acc[2], acc[3])));
  tc5::tmem_wait_st();
}

// O row (fp32, TMEM) * inv -> 16-bit row in global memory.
template <typename T, int D>
__device__ __forceinline__ void store_row(uint32_t o_addr, T* out, float inv, bool valid) {
  using M = MmaType<T>;
#pragma unroll
  for (int c = 0; c < D / 32; ++c) {
    uint32_t o[32];
    tc5::tmem_ld32(o_addr + c * 32, o);
    tc5::tmem_wait_ld();
    if (valid) {
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
}

// Which keys a CTA's queries see.  Dense (mvl_flash_prefill): a sequence's n new tokens attend to
// each other, query i to keys 0..i.  Paged (mvl_flash_prefill_paged): the sequence has ctx >= n keys
// in the paged cache (a cached prefix of ctx - n tokens, then the n new ones, already stored); query i
// sits at position ctx - n + i and sees keys 0..ctx - n + i.
struct SeqGeom {
  int q_start, n_new, ctx, prefix;
};
template <bool PAGED>
__device__ __forceinline__ SeqGeom seq_geom(const PrefillParams& p, int seq) {
  SeqGeom g;
  g.q_start = p.cu_seqlens[seq];
  g.n_new = p.cu_seqlens[seq + 1] - g.q_start;
  g.ctx = g.n_new;
  if (PAGED) {
    g.ctx = p.ctx_is_i64 ? (int)reinterpret_cast<const long long*>(p.context_lens)[seq]
                         : reinterpret_cast<const int*>(p.context_lens)[seq];
    if (g.ctx < g.n_new) g.ctx = g.n_new;
  }
  g.prefix = g.ctx - g.n_new;
  return g;
}
// Key tiles a 128-query tile starting at absolute position qa0 needs: through the tile of its last row.
__device__ __forceinline__ int key_tiles(int qa0, int ctx) { return min((qa0 + 127) / 128 + 1, (ctx + 127) >> 7); }

// TMA loads of key tile j (128 keys x D) into a 128-byte-swizzled shared-memory tile, issued
// cooperatively: lane li of the calling warp issues box li of the tile's D/64 panels x pieces.  Dense:
// one piece, 128 consecutive tokens of the packed tensor.  Paged: 128 / piece_rows pieces of
// piece_rows slots each, located through the block-table row (each lane reads its own entry); a piece
// past the context or with a negative table entry is requested at slot = num_slots, outside the
// tensor map, which the TMA unit fills with zeros (and still counts as bytes on the barrier).
// The caller posts expect_tx(TILE_BYTES) on the barrier first.
template <int D, bool PAGED>
__device__ __forceinline__ void kv_load(const PrefillParams& p, const SeqGeom& g, const int* table,
                                        const CUtensorMap* map, uint32_t dst, int head, int j, int li,
                                        uint32_t bar) {
  constexpr int PANELS = D / 64, PANEL_BYTES = 128 * 128;
  const int pieces = PAGED ? 128 / p.piece_rows : 1;
  if (li >= pieces * PANELS) return;
  const int pc = li / PANELS, pn = li % PANELS;
  int slot = g.q_start + j * 128, rows = 0;
  if (PAGED) {
    rows = pc * p.piece_rows;
    const int key0 = j * 128 + rows;
    slot = p.num_slots;
    if (key0 < g.ctx) {
      const int e = __ldg(table + (key0 >> p.bs_shift));
      if (e >= 0) slot = (e << p.bs_shift) | (key0 & ((1 << p.bs_shift) - 1));
    }
  }
  tma_load_3d(dst + pn * PANEL_BYTES + rows * 128, map, pn * 64, head, slot, bar);
}

template <typename T, int D, bool PAGED>
__global__ void __launch_bounds__(128, 2)
flash_prefill_tcgen05_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                             const __grid_constant__ CUtensorMap mv, const PrefillParams p) {
  using C = Tc5Cfg<D>;
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
  const SeqGeom g = seq_geom<PAGED>(p, seq);
  const int m0 = mb * C::BM;
  if (m0 >= g.n_new) return;
  const int qa0 = g.prefix + m0;              // absolute position of the tile's first query
  const int ntiles = key_tiles(qa0, g.ctx);   // the last one or two hold the causal diagonal
  const int* table = PAGED ? p.block_tables + (long long)seq * p.bt_stride : nullptr;

  if (tid == 64) tc5::prefetch_tensormap(&mq), tc5::prefetch_tensormap(&mk), tc5::prefetch_tensormap(&mv);
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

  // warp 0 issues the TMA boxes of a tile together (up to 16 of them for 16-slot pages)
  auto load_kv = [&](const CUtensorMap* map, uint32_t dst, int j, uint32_t bar) {
    if (tid == 0) mbar_expect_tx(bar, C::TILE_BYTES);
    __syncwarp();
    kv_load<D, PAGED>(p, g, table, map, dst, kvh, j, tid, bar);
  };
  constexpr uint32_t HI = tc5::smem_desc_hi_sw128(1024);
  const uint32_t q_lo = tc5::smem_desc_lo(sQ, 16), k_lo = tc5::smem_desc_lo(sK, 16),
                 v_lo = tc5::smem_desc_lo(sV, C::PANEL_BYTES);
  auto issue_qk = [&]() {  // S = Q . K^T: D/16 steps of K = 16 (32 bytes along the swizzled row)
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      const uint32_t off = ((ks >> 2) * C::PANEL_BYTES + (ks & 3) * 32) >> 4;
      tc5::mma_ss(tmem, tc5::desc64(q_lo + off, HI), tc5::desc64(k_lo + off, HI), IDESC_QK, ks > 0);
    }
    tc5::commit(bar_s);
  };
  auto issue_pv = [&](int j) {  // O += P . V: 8 steps of 16 keys (8 TMEM columns of P, 16 rows of V)
#pragma unroll
    for (int kk = 0; kk < C::BN / 16; ++kk)
      tc5::mma_ts(tmem + C::O_COL, tmem + kk * 8, tc5::desc64(v_lo + kk * (2048 >> 4), HI), IDESC_PV,
                  (j > 0 || kk > 0) ? 1u : 0u);
    tc5::commit(bar_o);
  };

  if (warp == 0) {
    if (tid == 0) {
      mbar_expect_tx(bar_q, C::TILE_BYTES);
#pragma unroll
      for (int pn = 0; pn < C::PANELS; ++pn)
        tma_load_3d(sQ + pn * C::PANEL_BYTES, &mq, pn * 64, h, g.q_start + m0, bar_q);
    }
    load_kv(&mk, sK, 0, bar_k);
    load_kv(&mv, sV, 0, bar_v);
    if (tid == 0) {
      tc5::mbar_wait_bounded(bar_q, 0);
      tc5::mbar_wait_bounded(bar_k, 0);
      tc5::fence_after_sync();
      issue_qk();
    }
    __syncwarp();
  }

  float m_ref = 0.f, l = 0.f;
  const float qs = p.qscale;
  for (int j = 0; j < ntiles; ++j) {
    tc5::mbar_wait_bounded(bar_s, j & 1);  // S_j is complete, and so is every MMA issued before it
    tc5::fence_after_sync();
    if (warp == 0) {
      if (j > 0) load_kv(&mv, sV, j, bar_v);                // product j-1 has retired
      if (j + 1 < ntiles) load_kv(&mk, sK, j + 1, bar_k);   // S_j no longer reads K_j
    }
    softmax_step<T, D>(lane_base, lane_base + C::O_COL, j == 0, qa0 + tid - j * 128, qs, m_ref, l);
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
  store_row<T, D>(lane_base + C::O_COL, reinterpret_cast<T*>(p.out) + ((long long)(g.q_start + row) * p.H + h) * D,
                  1.f / l, row < g.n_new);
  tc5::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc5::tmem_dealloc(tmem, C::TMEM_COLS);
}


// ---------------------------------------------------------------------------------------------
// Pair kernel: one CTA = 256 consecutive queries (two 128-row tiles) of one sequence for one head,
// one CTA per SM.  Warps 0-3 run the softmax of tile 0, warps 4-7 of tile 1; lane 0 of warp 8 issues
// every MMA.  Each tile has its own S (and aliased P) and O in tensor memory, so
// the issue order   P0V(j), QK0(j+1), P1V(j), QK1(j+1), ...   keeps the tensor pipe busy with one
// tile's products while the other tile's softmax runs.  K and V are double-buffered and shared by
// the two tiles; lane 0 of warp 9 issues the TMA loads, waiting on "empty" barriers that the MMA
// thread arms with tcgen05.commit after the last product reading a buffer.
template <int D>
struct PairCfg {
  static constexpr int PANELS = D / 64, PANEL_BYTES = 128 * 128, TILE_BYTES = PANELS * PANEL_BYTES;
  static constexpr int OFF_K = 2 * TILE_BYTES, OFF_V = 4 * TILE_BYTES, OFF_BAR = 6 * TILE_BYTES;
  static constexpr int NEED = OFF_BAR + 160 + 1024;
  // one CTA per SM (it allocates all 512 TMEM columns): ask for more than half the SM's shared memory
  static constexpr int SMEM = NEED > 120 * 1024 ? NEED : 120 * 1024;
  static constexpr int TMEM_COLS = 512;  // S0 [0,128) S1 [128,256) O0 [256,256+D) O1 [256+D, 256+2D)
};

template <typename T, int D, bool PAGED>
__global__ void __launch_bounds__(384, 1)
flash_prefill_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                                  const __grid_constant__ CUtensorMap mv, const PrefillParams p) {
  using C = PairCfg<D>;
  constexpr int FMT = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
  constexpr uint32_t IDESC_QK = tc5::instr_desc_f16(FMT, 0, 128, 128);
  constexpr uint32_t IDESC_PV = tc5::instr_desc_f16(FMT, 1, 128, D);
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sK = base + C::OFF_K, sV = base + C::OFF_V, bars = base + C::OFF_BAR;
  const uint32_t q_full = bars, k_full = bars + 8, v_full = bars + 24, s_full = bars + 40, p_ready = bars + 56,
                 o_done = bars + 72, k_empty = bars + 88, v_empty = bars + 104, p_half = bars + 120,
                 tmem_slot = bars + 136;  // pairs of barriers: [i] at + 8 * i

  const int tid = threadIdx.x, warp = tid >> 5;
  const int mb2 = gridDim.x - 1 - blockIdx.x;  // heaviest (latest) query tiles first
  const int h = blockIdx.y, seq = blockIdx.z;
  const int kvh = h / (p.H / p.KVH);
  const SeqGeom g = seq_geom<PAGED>(p, seq);
  const int m0 = mb2 * 256;
  if (m0 >= g.n_new) return;
  const int qa0 = g.prefix + m0;                                        // absolute position of query tile 0
  const int n0 = key_tiles(qa0, g.ctx);                                 // key tiles of query tile 0
  const int n1 = m0 + 128 < g.n_new ? key_tiles(qa0 + 128, g.ctx) : 0;  // query tile 1, if the sequence reaches it
  const int n = n1 ? n1 : n0;                                           // n1 >= n0 when tile 1 is live

  if (tid == 288) tc5::prefetch_tensormap(&mq), tc5::prefetch_tensormap(&mk), tc5::prefetch_tensormap(&mv);
  if (warp == 8) {
    tc5::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc5::tmem_relinquish();
  }
  if (tid == 0) {
    mbar_init(q_full, 1);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + 8 * i, 1), mbar_init(v_full + 8 * i, 1), mbar_init(s_full + 8 * i, 1);
      mbar_init(p_ready + 8 * i, 128), mbar_init(o_done + 8 * i, 1);
      mbar_init(k_empty + 8 * i, 1), mbar_init(v_empty + 8 * i, 1), mbar_init(p_half + 8 * i, 128);
    }
    mbar_fence_init();
  }
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = lds_u32(tmem_slot);

  if (warp >= 8) {
    // issue warpgroup: hands most of its registers to the two softmax warpgroups
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (warp == 9) {
      // ---- TMA warp: Q once, then K_j / V_j into stage j & 1 as soon as the products that read the
      // previous occupant (tile j - 2) have retired (k_empty / v_empty, signalled by tcgen05.commit).
      // The boxes of a tile (2 dense, up to 16 for 16-slot pages) are issued by as many lanes at once.
      const int lane = tid & 31;
      const int* table = PAGED ? p.block_tables + (long long)seq * p.bt_stride : nullptr;
      if (lane == 0) {
        mbar_expect_tx(q_full, (n1 ? 2 : 1) * C::TILE_BYTES);
#pragma unroll
        for (int pn = 0; pn < C::PANELS; ++pn) {
          tma_load_3d(sQ + pn * C::PANEL_BYTES, &mq, pn * 64, h, g.q_start + m0, q_full);
          if (n1)
            tma_load_3d(sQ + C::TILE_BYTES + pn * C::PANEL_BYTES, &mq, pn * 64, h, g.q_start + m0 + 128, q_full);
        }
      }
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        if (j >= 2) tc5::mbar_wait_bounded(k_empty + 8 * st, ((j >> 1) - 1) & 1);
        if (lane == 0) mbar_expect_tx(k_full + 8 * st, C::TILE_BYTES);
        __syncwarp();
        kv_load<D, PAGED>(p, g, table, &mk, sK + st * C::TILE_BYTES, kvh, j, lane, k_full + 8 * st);
        if (j >= 2) tc5::mbar_wait_bounded(v_empty + 8 * st, ((j >> 1) - 1) & 1);
        if (lane == 0) mbar_expect_tx(v_full + 8 * st, C::TILE_BYTES);
        __syncwarp();
        kv_load<D, PAGED>(p, g, table, &mv, sV + st * C::TILE_BYTES, kvh, j, lane, v_full + 8 * st);
      }
    } else if (tid == 256) {
      // ---- MMA thread
      constexpr uint32_t HI = tc5::smem_desc_hi_sw128(1024);
      constexpr uint32_t STAGE = C::TILE_BYTES >> 4;
      const uint32_t q_lo = tc5::smem_desc_lo(sQ, 16), k_lo = tc5::smem_desc_lo(sK, 16),
                     v_lo = tc5::smem_desc_lo(sV, C::PANEL_BYTES);
      auto issue_qk = [&](int w, int j) {  // S_w = Q_w . K_j^T
        const uint32_t q = q_lo + w * STAGE, k = k_lo + (j & 1) * STAGE;
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const uint32_t off = ((ks >> 2) * C::PANEL_BYTES + (ks & 3) * 32) >> 4;
          tc5::mma_ss(tmem + w * 128, tc5::desc64(q + off, HI), tc5::desc64(k + off, HI), IDESC_QK, ks > 0);
        }
        tc5::commit(s_full + 8 * w);
      };
      auto issue_pv = [&](int w, int j) {  // O_w += P_w . V_j, the first 64 keys as soon as their P is in TMEM
        const uint32_t v = v_lo + (j & 1) * STAGE;
        tc5::mbar_wait_bounded(p_half + 8 * w, j & 1);
        tc5::fence_after_sync();
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          tc5::mma_ts(tmem + 256 + w * D, tmem + w * 128 + kk * 8, tc5::desc64(v + kk * (2048 >> 4), HI), IDESC_PV,
                      (j > 0 || kk > 0) ? 1u : 0u);
        tc5::mbar_wait_bounded(p_ready + 8 * w, j & 1);
        tc5::fence_after_sync();
#pragma unroll
        for (int kk = 4; kk < 8; ++kk)
          tc5::mma_ts(tmem + 256 + w * D, tmem + w * 128 + kk * 8, tc5::desc64(v + kk * (2048 >> 4), HI), IDESC_PV, 1u);
      };
      tc5::mbar_wait_bounded(q_full, 0);
      tc5::mbar_wait_bounded(k_full, 0);
      tc5::fence_after_sync();
      issue_qk(0, 0);
      if (n1) issue_qk(1, 0);
      tc5::commit(k_empty);
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        tc5::mbar_wait_bounded(v_full + 8 * st, (j >> 1) & 1);
        bool k_next_seen = false;
        if (j < n0) {
          issue_pv(0, j);
          if (j == n0 - 1) tc5::commit(o_done);
          if (j >= n1) tc5::commit(v_empty + 8 * st);   // tile 1 does not read V_j
          if (j + 1 < n0) {
            tc5::mbar_wait_bounded(k_full + 8 * (st ^ 1), ((j + 1) >> 1) & 1);
            tc5::fence_after_sync();
            k_next_seen = true;
            issue_qk(0, j + 1);
            if (j + 1 >= n1) tc5::commit(k_empty + 8 * (st ^ 1));
          }
        }
        if (j < n1) {
          issue_pv(1, j);
          if (j == n1 - 1) tc5::commit(o_done + 8);
          tc5::commit(v_empty + 8 * st);
          if (j + 1 < n1) {
            if (!k_next_seen) {
              tc5::mbar_wait_bounded(k_full + 8 * (st ^ 1), ((j + 1) >> 1) & 1);
              tc5::fence_after_sync();
            }
            issue_qk(1, j + 1);
            tc5::commit(k_empty + 8 * (st ^ 1));
          }
        }
      }
    }
    __syncwarp();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    const int w = warp >> 2, t = tid & 127;
    const int nw = w ? n1 : n0;
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t s_addr = lane_base + w * 128, o_addr = lane_base + 256 + w * D;
    float m_ref = 0.f, l = 0.f;
    const float qs = p.qscale;
    for (int j = 0; j < nw; ++j) {
      tc5::mbar_wait_bounded(s_full + 8 * w, j & 1);  // S_w(j) complete; so is the product of step j-1
      tc5::fence_after_sync();
      softmax_step<T, D>(s_addr, o_addr, j == 0, qa0 + w * 128 + t - j * 128, qs, m_ref, l, p_half + 8 * w);
      tc5::fence_before_sync();
      tc5::mbar_arrive(p_ready + 8 * w);
    }
    if (nw > 0) {
      tc5::mbar_wait_bounded(o_done + 8 * w, 0);
      tc5::fence_after_sync();
      const int row = m0 + w * 128 + t;
      store_row<T, D>(o_addr, reinterpret_cast<T*>(p.out) + ((long long)(g.q_start + row) * p.H + h) * D, 1.f / l,
                      row < g.n_new);
    }
  }
  tc5::fence_before_sync();
  __syncthreads();
  if (warp == 8) tc5::tmem_dealloc(tmem, C::TMEM_COLS);
}

template <typename T, int D, bool PAGED>
int launch_pair_td(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const PrefillParams& p,
                   int num_seqs, int max_seqlen, cudaStream_t s) {
  using C = PairCfg<D>;
  auto kern = flash_prefill_tcgen05_pair_kernel<T, D, PAGED>;
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) configured[dev] = true;
  }
  dim3 grid((unsigned)((max_seqlen + 255) / 256), (unsigned)p.H, (unsigned)num_seqs);
  kern<<<grid, 384, C::SMEM, s>>>(mq, mk, mv, p);
  return (int)cudaGetLastError();
}

template <typename T, int D, bool PAGED>
int launch_td(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const PrefillParams& p,
              int num_seqs, int max_seqlen, cudaStream_t s) {
  using C = Tc5Cfg<D>;
  auto kern = flash_prefill_tcgen05_kernel<T, D, PAGED>;
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

template <typename T, int D>
int launch_any(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const PrefillParams& p,
               int num_seqs, int max_seqlen, bool single_tile, cudaStream_t s) {
  const bool paged = p.block_tables != nullptr;
  if (single_tile)
    return paged ? launch_td<T, D, true>(mq, mk, mv, p, num_seqs, max_seqlen, s)
                 : launch_td<T, D, false>(mq, mk, mv, p, num_seqs, max_seqlen, s);
  return paged ? launch_pair_td<T, D, true>(mq, mk, mv, p, num_seqs, max_seqlen, s)
               : launch_pair_td<T, D, false>(mq, mk, mv, p, num_seqs, max_seqlen, s);
}

}  // namespace

int launch_prefill_tcgen05(const PrefillParams& p, int dtype, int head_dim, int num_seqs, int max_seqlen,
                           bool single_tile, cudaStream_t s) {
  const bool paged = p.block_tables != nullptr;
  // K / V source: the packed (T, KVH, D) tensors, or the paged caches (num_slots, KVH, D) in boxes of
  // piece_rows slots
  const long long kv_rows = paged ? p.num_slots : p.total_tokens;
  const int kv_box = paged ? p.piece_rows : 128;
  CUtensorMap mq, mk, mv;
  int r;
  if ((r = make_packed_map(&mq, p.q, dtype, head_dim, p.H, p.total_tokens, p.q_stride, 128)) != 0)
    return fail_text("prefill tensor map (q)", r);
  if ((r = make_packed_map(&mk, p.k, dtype, head_dim, p.KVH, kv_rows, p.k_stride, kv_box)) != 0)
    return fail_text("prefill tensor map (k)", r);
  if ((r = make_packed_map(&mv, p.v, dtype, head_dim, p.KVH, kv_rows, p.v_stride, kv_box)) != 0)
    return fail_text("prefill tensor map (v)", r);
  if (dtype == 2)
    r = head_dim == 64 ? launch_any<__nv_bfloat16, 64>(mq, mk, mv, p, num_seqs, max_seqlen, single_tile, s)
                       : launch_any<__nv_bfloat16, 128>(mq, mk, mv, p, num_seqs, max_seqlen, single_tile, s);
  else
    r = head_dim == 64 ? launch_any<__half, 64>(mq, mk, mv, p, num_seqs, max_seqlen, single_tile, s)
                       : launch_any<__half, 128>(mq, mk, mv, p, num_seqs, max_seqlen, single_tile, s);
  return r == 0 ? 0 : fail_cuda(r, "flash_prefill_tcgen05 launch");
}

}  // namespace mvl
