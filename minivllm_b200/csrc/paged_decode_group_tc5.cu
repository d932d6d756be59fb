// paged_decode_group_tc5.cu — prefix sharing, wide form, on the 5th-generation tensor cores.
//
// What it computes: for a chunk of tokens that all members of a sub-group of
// sequences share (their block tables alias: the reference's prefix-caching block manager,
// src/myvllm/engine/block_manager.py:39-44, 67-97), the un-normalised attention partial of EVERY
// (member, query head) pair of one kv head against ONE copy of the chunk's K/V — the same partials the
// split kernel (paged_decode_mma.cuh) would have written sequence by sequence, in its layout, so that
// its fused combine merges them with the members' private chunks.
//
// Why tcgen05.  The mma.sync group kernel (paged_decode_group.cuh) gives each of its four warps a
// quarter of the members and lets all four walk every tile: a tile crosses L2 -> SM once, but costs four
// warps' worth of ldmatrix + mma.sync + softmax per byte, and C5's shared chunks run issue-bound at about
// half the HBM rate.  Here the sub-group's (member, head) pairs — at most 32 — are the ROWS of one
// M = 128 accumulator tile, a 128-token K/V tile is the N (scores) and then the K (P.V) dimension:
//   S (128 x 128 fp32, TMEM) = Q (rows x D, smem)  . K_tile^T (128 tokens x D, smem)     D/16 tcgen05.mma
//   O (128 x D  fp32, TMEM) += P (rows x 128, TMEM) . V_tile   (128 tokens x D, smem)     8 tcgen05.mma
// One thread issues all of it; rows 32..127 of the tile are padding the tensor pipe does not miss.
//
// Shape of the kernel: one CTA per SM, persistent over the work items (sub-group, kv head, shared chunk)
// that the device pre-pass listed (prefix_share.cu), 128 threads in four roles:
//   warp 0  softmax: thread r owns row r = TMEM lane r (member r >> log2 Gp, head r & (Gp - 1)); a score
//           row arrives with tcgen05.ld, P goes back over the first 64 columns of S as 16-bit pairs
//   warps 1, 3  K/V producers: a 3-stage ring of (K tile, V tile) = 3 x 64 KB filled with 16-byte cp.async
//           copies, rows located through the leader's block-table row (the members' rows agree on these
//           pages), written in the 128-byte-swizzled layout the MMA descriptors read; "full" barriers
//           by cp.async.mbarrier.arrive, "empty" barriers armed by tcgen05.commit.  (Q comes by TMA.)
//   warp 2  lane 0 issues every MMA, in the order QK(j+1), PV(j): the next tile's scores are computed
//           while this tile's softmax runs (S is double-buffered in TMEM)
// With 192 KB of K/V in flight per SM the loads are never exposed (the first tcgen05 attempt of round 1
// had one K and one V buffer per CTA and lost to the mma.sync kernel for that reason).
//
// Launched in FRONT of the split kernel, which follows as its programmatic dependent (the two touch
// disjoint chunks): the 148 big CTAs start at once, the split kernel's small ones fill an SM when its
// group CTA leaves.
#include "common.cuh"
#include "host.h"
#include "paged_decode_mma.cuh"
#include "prefill.h"
#include "tcgen05.cuh"

namespace mvl {
namespace {

// TPC = tokens per K/V tile.  64 (six stages of 16 + 16 KB) rather than 128 (three of 32 + 32 KB): a stage
// is busy from the moment the producer gets it until the P.V of its tile retires — copy issue, DRAM
// latency, Q.K, softmax, P.V — and with three big stages that round trip (4.6 us measured) capped a CTA at
// one 128-token tile per 1.5 us, below what the SM's share of HBM delivers.
template <int D, int TPC_ = 64>
struct Gt5Cfg {
  static constexpr int TPC = TPC_;
  static constexpr int STAGES = 384 / TPC;          // 192 KB of K/V in flight per SM at D = 128
  static constexpr int PANELS = D / 64;             // 64 elements = one 128-byte swizzle row
  static constexpr int Q_PANEL_BYTES = 128 * 128;   // Q tile: 128 rows x 128 bytes per panel
  static constexpr int Q_TILE_BYTES = PANELS * Q_PANEL_BYTES;
  static constexpr int PANEL_BYTES = TPC * 128;     // K / V tile: TPC rows x 128 bytes per panel
  static constexpr int TILE_BYTES = PANELS * PANEL_BYTES;
  static constexpr int OFF_KV = Q_TILE_BYTES;       // stage s: K at OFF_KV + 2 s TILE_BYTES, V one tile further
  static constexpr int OFF_BAR = OFF_KV + 2 * STAGES * TILE_BYTES;
  static constexpr int SMEM = OFF_BAR + 256 + 1024;  // + barriers / TMEM slot / flags + alignment slack
  // S0 [0, TPC)  S1 [TPC, 2 TPC)  O of the even tiles [2 TPC, 2 TPC + D)  O of the odd tiles behind it
  static constexpr int O_COL = 2 * TPC;
  static constexpr int TMEM_COLS = (2 * TPC + 2 * D) <= 256 ? 256 : 512;
};

__device__ __forceinline__ float fmax3f(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// One softmax step of one row (thread = row = TMEM lane) over a 128-token tile, every token visible.
// Lazy maximum: O is rescaled in TMEM only when the row maximum grew by more than 2^8 (log2 units);
// before touching O the product of the previous tile must have retired (bar_o, phase o_phase).
template <typename T, int D, int TPC>
__device__ __forceinline__ void group_softmax_step(uint32_t s_addr, uint32_t o_addr, bool first, float qs,
                                                   float& m_ref, float& l, uint32_t bar_o, uint32_t o_phase) {
  using M = MmaType<T>;
  uint32_t su[TPC];
#pragma unroll
  for (int c = 0; c < TPC / 32; ++c) tc5::tmem_ld32(s_addr + c * 32, su + c * 32);
  tc5::tmem_wait_ld();
  float mxa[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int c = 0; c < TPC; c += 8)
#pragma unroll
    for (int k = 0; k < 4; ++k)
      mxa[k] = fmax3f(mxa[k], __uint_as_float(su[c + 2 * k]), __uint_as_float(su[c + 2 * k + 1]));
  const float mx = fmaxf(fmaxf(mxa[0], mxa[1]), fmaxf(mxa[2], mxa[3]));
  bool grow = false;
  if (first) m_ref = mx;
  else grow = (mx - m_ref) * qs > 8.f;
  if (__any_sync(0xffffffffu, grow)) {  // warp-uniform: the TMEM moves are warp-collective
    tc5::mbar_wait_bounded(bar_o, o_phase);  // the previous tile's P.V no longer adds into O
    tc5::fence_after_sync();
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
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int part = 0; part < TPC / 64; ++part) {   // 64 scores -> 32 columns of 16-bit pairs
    uint32_t pk[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      const int cc = part * 32 + c;
      const float p0 = fast_exp2(fmaf(__uint_as_float(su[2 * cc]), qs, neg));
      const float p1 = fast_exp2(fmaf(__uint_as_float(su[2 * cc + 1]), qs, neg));
      pk[c] = M::pack(p0, p1);
      // the sums use what the products will use: p rounded to 16 bits
      acc[c & 3] += M::lo_f(pk[c]) + M::hi_f(pk[c]);
    }
    tc5::tmem_st32(s_addr + part * 32, pk);
  }
  l += (acc[0] + acc[1]) + (acc[2] + acc[3]);
  tc5::tmem_wait_st();
}

__device__ __forceinline__ uint32_t lds_word(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

// (4 stages and a 184-register cap leave room on every SM for one CTA of the split kernel beside this one;
// measured on C5: items 52 -> 59 us, step 0.351 -> 0.359 ms — sharing the SM costs this kernel more than the
// split kernel's single CTA gains.  One CTA per SM, all of it.)
template <typename T, int D, bool PAGE16>
__global__ void __launch_bounds__(256, 1)
paged_decode_group_tc5_kernel(const __grid_constant__ CUtensorMap mq, const DecodeParams p, const int batch) {
  using C = Gt5Cfg<D>;
  constexpr int FMT = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
  constexpr int STAGES = C::STAGES, TPC = C::TPC;  // tokens per tile
  constexpr uint32_t IDESC_QK = tc5::instr_desc_f16(FMT, 0, 128, TPC);
  constexpr uint32_t IDESC_PV = tc5::instr_desc_f16(FMT, 1, 128, D);
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;
  auto sK = [&](int st) -> uint32_t { return base + C::OFF_KV + (uint32_t)st * 2 * C::TILE_BYTES; };
  auto sV = [&](int st) -> uint32_t { return sK(st) + C::TILE_BYTES; };
  const uint32_t bar0 = base + C::OFF_BAR;
  const uint32_t bar_q = bar0;
  auto bar_kf = [&](int st) { return bar0 + 8 + 8 * st; };                      // K tile of a stage landed
  auto bar_vf = [&](int st) { return bar0 + 8 + 8 * (STAGES + st); };           // V tile landed
  auto bar_emp = [&](int st) { return bar0 + 8 + 8 * (2 * STAGES + st); };      // both consumed
  auto bar_sf = [&](int sb) { return bar0 + 8 + 8 * (3 * STAGES + sb); };       // S buffer holds a tile's scores
  auto bar_pr = [&](int sb) { return bar0 + 8 + 8 * (3 * STAGES + 2 + sb); };   // ... its P
  auto bar_o = [&](int sb) { return bar0 + 8 + 8 * (3 * STAGES + 4 + sb); };    // P.V into O[sb] retired
  const uint32_t tmem_slot = bar0 + 8 + 8 * (3 * STAGES + 6), flag = tmem_slot + 4;
  static_assert(8 + 8 * (3 * STAGES + 6) + 8 <= 256, "barrier block");

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 64) tc5::prefetch_tensormap(&mq);
  if (warp == 2) {
    tc5::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc5::tmem_relinquish();
  }
  if (tid == 32) {
    mbar_init(bar_q, 1);
    for (int st = 0; st < STAGES; ++st) mbar_init(bar_kf(st), 64), mbar_init(bar_vf(st), 64), mbar_init(bar_emp(st), 1);
    for (int sb = 0; sb < 2; ++sb) mbar_init(bar_sf(sb), 1), mbar_init(bar_pr(sb), 1);
    mbar_init(bar_o(0), 1), mbar_init(bar_o(1), 1);
    mbar_fence_init();
  }
  // rows of the Q tile the loads never write (rows past the sub-group's 32) must hold finite garbage at most
  for (uint32_t i = tid * 16; i < (uint32_t)C::Q_TILE_BYTES; i += 256 * 16)
    asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(sQ + i), "r"(0u) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = lds_word(tmem_slot);

  // Everything above ran while the pre-pass in front (programmatic dependency) was still finishing.  Its
  // work list is read only behind this wait — and only then may the split kernel, OUR dependent, start:
  // it reads the same pre-pass output and does not wait at its entry.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // ---- work list of the pre-pass: wl_b[e] = leader of wide sub-group e, wl_cum[e] = shared chunks of
  // the sub-groups before it; meta = leader | gsize | gshare | lcpt | wl_b | wl_cum | (n, total)
  const int* meta = p.share_meta;
  const int* gsize = meta + batch;
  const int* wl_b = meta + 4 * batch;
  const int* wl_cum = meta + 5 * batch;
  const int wl_n = meta[6 * batch], wl_total = meta[6 * batch + 1];
  const int X = p.KVH;                       // kv heads (sharing needs G <= 4: one head tile each)
  const long long n_items = (long long)wl_total * X;
  // as many CTAs as make every one of them do the same number of items (416 items on 148 SMs: 139 CTAs x 3
  // rather than 120 x 3 + 28 x 2); the SMs left over go to the split kernel from the start
  const long long rounds = n_items > 0 ? (n_items + gridDim.x - 1) / gridDim.x : 1;
  const int use_ctas = (int)((n_items + rounds - 1) / rounds);
  if ((int)blockIdx.x >= use_ctas) {  // nothing for this CTA (use_ctas = 0 when no sub-group is wide)
    __syncthreads();
    if (warp == 2) tc5::tmem_dealloc(tmem, C::TMEM_COLS);
    return;
  }

  // diagnostics (mvl_debug_set_times): CTA 0's first item, 6 stamps per tile
  auto stamp = [&](int j, int which) {
    if (p.dbg_times != nullptr && blockIdx.x == 0 && j < 32) {
      unsigned long long tnow;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tnow));
      p.dbg_times[j * 6 + which] = tnow;
    }
  };
  const int gp = 1 << p.gp_shift, gp_mask = gp - 1;
  const int rows_box = 32;                          // (32 / gp) members x gp heads
  const int tiles = p.chunk / TPC;                  // a shared chunk is a whole chunk (multiple of 128)
  const uint32_t q_bytes = (uint32_t)C::PANELS * rows_box * 128;
  constexpr uint32_t HI = tc5::smem_desc_hi_sw128(1024);
  const uint32_t q_lo = tc5::smem_desc_lo(sQ, 16);

  uint32_t gt = 0;      // tiles this CTA has streamed so far (all roles count alike): stage / phase bookkeeping
  uint32_t items_done = 0;
  for (long long item = blockIdx.x; item < n_items; item += use_ctas, ++items_done) {
    // ---- which (sub-group, kv head, chunk)
    const int x = (int)(item % X);
    const int ci = (int)(item / X);
    int lo = 0, hi = wl_n;  // wl_cum[lo] <= ci < wl_cum[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (wl_cum[mid] <= ci) lo = mid; else hi = mid;
    }
    const int b = wl_b[lo], split = ci - wl_cum[lo];
    const int nmem = gsize[b];
    const int start = split * p.chunk;
    const int head0 = x * p.G;
    const int g_valid = p.G;
    const uint32_t q_phase = items_done & 1u;
    auto stamp_item = [&](int which) {   // diagnostics: per CTA and item, start / streamed / counted / merged
      if (p.dbg_times != nullptr && tid == 0 && items_done < 4) {
        unsigned long long tnow;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tnow));
        p.dbg_times[192 + (blockIdx.x * 4 + items_done) * 5 + which] = tnow;
      }
    };
    stamp_item(0);

    if (warp == 1 || warp == 3) {
      // ================================================================ K/V producer (64 threads)
      // 16-byte cp.async copies straight into the 128-byte-swizzled tile layout the MMA descriptors
      // expect (row = token, 128 bytes per row and panel, chunk index XOR row % 8).  Measured first with
      // TMA boxes: with 16-slot pages a tile is 32 boxes of 2 KB, and the TMA unit's box rate capped the
      // kernel at 2.2 TB/s.  A thread owns chunk column `cc` of rows rg, rg + 4, ...; it arrives on the
      // stage's barriers with cp.async.mbarrier.arrive.noinc, i.e. when its copies have landed.
      const int pt = (warp == 1 ? 0 : 32) + lane;        // 0..63
      const int cc = pt % (C::PANELS * 8), rg = pt / (C::PANELS * 8);
      constexpr int RSTEP = 64 / (C::PANELS * 8);         // rows covered per pass of the 64 threads
      if (pt == 0) {
        mbar_expect_tx(bar_q, q_bytes);
#pragma unroll
        for (int pn = 0; pn < C::PANELS; ++pn) tma_load_3d(sQ + pn * C::Q_PANEL_BYTES, &mq, pn * 64, head0, b, bar_q);
      }
      const int32_t* table = p.bt + (long long)b * p.bt_stride;
      const int bs_mask = (1 << p.bs_shift) - 1;
      const uint32_t row_pitch = (uint32_t)p.KVH * (D * 2);
      const char* kbase = reinterpret_cast<const char*>(p.kc) + (long long)x * (D * 2) + cc * 16;
      const char* vbase = reinterpret_cast<const char*>(p.vc) + (long long)x * (D * 2) + cc * 16;
      const uint32_t col_off = (uint32_t)(cc >> 3) * C::PANEL_BYTES;
      // row r = rg + kk (kk a multiple of RSTEP): swizzled offset inside its 8-row atom, and byte offset of
      // its slot inside a 16-slot page, for the few values (kk & 7) and (kk & 15) can take
      uint32_t rowoff[8 / RSTEP];
      long long within[16 / RSTEP];
#pragma unroll
      for (int u = 0; u < 8 / RSTEP; ++u) {
        const int r7 = rg + u * RSTEP;
        rowoff[u] = (uint32_t)r7 * 128u + (uint32_t)(((cc & 7) ^ r7) << 4);
      }
#pragma unroll
      for (int u = 0; u < 16 / RSTEP; ++u) within[u] = (long long)(rg + u * RSTEP) * row_pitch;
      for (int j = 0; j < tiles; ++j) {
        const uint32_t g = gt + j;
        const int st = g % STAGES;
        if (g >= (uint32_t)STAGES) tc5::mbar_wait_bounded(bar_emp(st), ((g / STAGES) - 1) & 1u);
        if (pt == 0 && items_done == 0) stamp(j, 0);
        const uint32_t kdst = sK(st) + col_off, vdst = sV(st) + col_off;
        const int tok0 = start + j * TPC;
        // the tile's table entries (at most 8 pages of >= 16 slots): lane l holds entry l, rows fetch
        // theirs with a shuffle — one table load per tile and lane instead of a dependent load per row
        const int pg0 = tok0 >> p.bs_shift;
        const int npg = ((tok0 + TPC - 1) >> p.bs_shift) - pg0 + 1;
        const uint32_t my_e = lane < npg ? (uint32_t)__ldg(table + pg0 + lane) : 0u;
        if (PAGE16) {
          // 16-slot pages (the reference's default, attention.py:483): page and slot-in-page of every row
          // this thread copies are compile-time functions of the unrolled row counter, so a copy costs an
          // add and the cp.async.  A producer warp has its scheduler to itself — nothing hides the
          // latency of a longer dependent address chain (measured: 23 instructions per copy, 4.4 us per tile).
          const char* kpage[TPC / 16];
#pragma unroll
          for (int pg = 0; pg < TPC / 16; ++pg)
            kpage[pg] = kbase + (unsigned long long)(__shfl_sync(0xffffffffu, my_e, pg) << 4) * row_pitch;
          const long long v_minus_k = vbase - kbase;
#pragma unroll
          for (int i = 0; i < TPC / RSTEP; ++i) {
            const int kk = RSTEP * i;  // r = rg + kk
            const uint32_t off = (uint32_t)(kk >> 3) * 1024u + rowoff[(kk & 7) / RSTEP];
            cp_async16(kdst + off, kpage[kk >> 4] + within[(kk & 15) / RSTEP], 16u);
          }
          asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar_kf(st)) : "memory");
#pragma unroll
          for (int i = 0; i < TPC / RSTEP; ++i) {
            const int kk = RSTEP * i;
            const uint32_t off = (uint32_t)(kk >> 3) * 1024u + rowoff[(kk & 7) / RSTEP];
            cp_async16(vdst + off, kpage[kk >> 4] + within[(kk & 15) / RSTEP] + v_minus_k, 16u);
          }
        } else {
#pragma unroll 8
          for (int r = rg; r < TPC; r += RSTEP) {
            const int tok = tok0 + r;
            const uint32_t e = __shfl_sync(0xffffffffu, my_e, (tok >> p.bs_shift) - pg0);
            const uint32_t slot = (e << p.bs_shift) + (uint32_t)(tok & bs_mask);
            const uint32_t off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + (uint32_t)(((cc & 7) ^ (r & 7)) << 4);
            cp_async16(kdst + off, row_ptr(kbase, slot, row_pitch), 16u);
          }
          asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar_kf(st)) : "memory");
#pragma unroll 8
          for (int r = rg; r < TPC; r += RSTEP) {
            const int tok = tok0 + r;
            const uint32_t e = __shfl_sync(0xffffffffu, my_e, (tok >> p.bs_shift) - pg0);
            const uint32_t slot = (e << p.bs_shift) + (uint32_t)(tok & bs_mask);
            const uint32_t off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + (uint32_t)(((cc & 7) ^ (r & 7)) << 4);
            cp_async16(vdst + off, row_ptr(vbase, slot, row_pitch), 16u);
          }
        }
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar_vf(st)) : "memory");
        if (pt == 0 && items_done == 0) stamp(j, 1);
      }
    } else if (warp == 2) {
      // ================================================================ MMA issuer
      if (lane == 0) {
        auto issue_qk = [&](uint32_t g) {
          const uint32_t k_lo = tc5::smem_desc_lo(sK(g % STAGES), 16);
          const uint32_t s_t = tmem + (g & 1u) * TPC;
#pragma unroll
          for (int ks = 0; ks < D / 16; ++ks) {
            const uint32_t qoff = ((ks >> 2) * C::Q_PANEL_BYTES + (ks & 3) * 32) >> 4;
            const uint32_t koff = ((ks >> 2) * C::PANEL_BYTES + (ks & 3) * 32) >> 4;
            tc5::mma_ss(s_t, tc5::desc64(q_lo + qoff, HI), tc5::desc64(k_lo + koff, HI), IDESC_QK, ks > 0);
          }
          tc5::commit(bar_sf(g & 1u));
        };
        // K/V tiles are written by cp.async (generic proxy) and read by the tensor cores (async proxy)
        auto proxy_fence = [] { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); };
        tc5::mbar_wait_bounded(bar_q, q_phase);
        tc5::mbar_wait_bounded(bar_kf(gt % STAGES), (gt / STAGES) & 1u);
        proxy_fence();
        tc5::fence_after_sync();
        issue_qk(gt);
        for (int j = 0; j < tiles; ++j) {
          const uint32_t g = gt + j;
          if (j + 1 < tiles) {
            tc5::mbar_wait_bounded(bar_kf((g + 1) % STAGES), ((g + 1) / STAGES) & 1u);
            proxy_fence();
            tc5::fence_after_sync();
            if (items_done == 0) stamp(j + 1, 2);
            issue_qk(g + 1);  // queued behind P.V of tile j-1, which read the S buffer this one overwrites
          }
          tc5::mbar_wait_bounded(bar_pr(g & 1u), (g >> 1) & 1u);
          tc5::mbar_wait_bounded(bar_vf(g % STAGES), (g / STAGES) & 1u);
          proxy_fence();
          tc5::fence_after_sync();
          if (items_done == 0) stamp(j, 3);
          const uint32_t v_lo = tc5::smem_desc_lo(sV(g % STAGES), C::PANEL_BYTES);
          const uint32_t p_t = tmem + (g & 1u) * TPC;
#pragma unroll
          for (int kk = 0; kk < TPC / 16; ++kk)
            tc5::mma_ts(tmem + C::O_COL + (j & 1) * D, p_t + kk * 8, tc5::desc64(v_lo + kk * (2048 >> 4), HI), IDESC_PV,
                        (j > 1 || kk > 0) ? 1u : 0u);
          tc5::commit(bar_emp(g % STAGES));
          tc5::commit(bar_o(j & 1));
        }
      }
      __syncwarp();
    } else if (warp == 0 || warp == 4) {
      // ================================================================ softmax: two streams
      // Warp 0 takes the even tiles, warp 4 the odd ones (both reach TMEM lanes 0..31), each with its own
      // running maximum, sum and O accumulator: one warp doing every tile's 64 x 32 exponentials was the
      // slowest link of the tile pipeline (measured: 0.8-1.0 us of softmax + 0.25 us hand-off per 128
      // tokens against 1.2 us of HBM time).  The two streams are merged below like two chunks.
      const int q = warp >> 2;
      float m_ref = 0.f, l = 0.f;
      const uint32_t lane_base = tmem;  // TMEM lanes 0..31
      const uint32_t o_addr = lane_base + C::O_COL + q * D;
      for (int j = q; j < tiles; j += 2) {
        const uint32_t g = gt + j;
        tc5::mbar_wait_bounded(bar_sf(q), (g >> 1) & 1u);
        tc5::fence_after_sync();
        if (lane == 0 && items_done == 0) stamp(j, 4);
        group_softmax_step<T, D, TPC>(lane_base + q * TPC, o_addr, j < 2, p.qscale, m_ref, l, bar_o(q),
                                      ((g >> 1) - 1) & 1u);
        tc5::fence_before_sync();
        __syncwarp();
        if (lane == 0 && items_done == 0) stamp(j, 5);
        if (lane == 0) tc5::mbar_arrive(bar_pr(q));
      }
      const uint32_t g_last = gt + tiles - 2 + q;
      tc5::mbar_wait_bounded(bar_o(q), (g_last >> 1) & 1u);
      tc5::fence_after_sync();
      // this stream's row: O (D fp32), m (log2 units), l -> registers; the odd stream parks its row in
      // shared memory (the K/V ring is idle: its wait above was for the item's last product)
      float orow[D];
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        tc5::tmem_ld32(o_addr + c * 32, o);
        tc5::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) orow[c * 32 + i] = __uint_as_float(o[i]);
      }
      const uint32_t xrow = sK(0) + (uint32_t)lane * (D + 4) * 4;
      if (q == 1) {
#pragma unroll
        for (int i = 0; i < D; i += 4)
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(xrow + i * 4), "f"(orow[i]), "f"(orow[i + 1]),
                       "f"(orow[i + 2]), "f"(orow[i + 3]) : "memory");
        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(xrow + D * 4), "f"(m_ref * p.qscale), "f"(l) : "memory");
      }
      tc5::fence_before_sync();
      // both softmax warps (64 threads) meet: named barrier 1
      asm volatile("bar.sync 1, 64;" ::: "memory");
      if (q == 0) {
        float mo, lo2;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(mo), "=f"(lo2) : "r"(xrow + D * 4) : "memory");
        const float me = m_ref * p.qscale;
        const float Mx = fmaxf(me, mo);
        const float se = fast_exp2(me - Mx), so = fast_exp2(mo - Mx);
        const float Lsum = l * se + lo2 * so;
        const int m = lane >> p.gp_shift, hd = lane & gp_mask;
        const bool row_ok = m < nmem && hd < g_valid;
        const long long pi = ((long long)(b + m) * p.H + head0 + hd) * p.max_splits + split;
#pragma unroll
        for (int i = 0; i < D; i += 4) {
          float x0, x1, x2, x3;
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x0), "=f"(x1), "=f"(x2), "=f"(x3) : "r"(xrow + i * 4) : "memory");
          if (row_ok)
            *reinterpret_cast<float4*>(p.part_o + pi * D + i) =
                make_float4(orow[i] * se + x0 * so, orow[i + 1] * se + x1 * so, orow[i + 2] * se + x2 * so,
                            orow[i + 3] * se + x3 * so);
        }
        if (row_ok) p.part_ml[pi] = make_float2(Mx, Lsum);
      }
    }
    gt += tiles;

    // ---- end of the item: every role is done with it (TMEM, the Q tile and the ring are free again);
    // then, per member, the last of its chunk CTAs to arrive merges its partials (split kernel's protocol)
    __syncthreads();
    tc5::fence_after_sync();
    stamp_item(1);
    auto member_chunks = [&](int mm) -> int {
      long long c = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b + mm]
                                 : (long long)reinterpret_cast<const int*>(p.ctx)[b + mm];
      if (c > p.max_ctx) c = p.max_ctx;
      return (int)((c + p.chunk - 1) / p.chunk);
    };
    if (p.counters != nullptr) {
      // lane m of warp 0 counts member m's arrival (the eight-odd members' atomics in parallel, not one
      // thread after the other: that loop cost 5 us per item)
      if (warp == 0) {
        __threadfence();
        bool last = false;
        if (lane < nmem) {
          const int na = member_chunks(lane);
          int* cnt = p.counters + (long long)(b + lane) * X + x;
          if (atomicAdd(cnt, 1) == na - 1) {
            *cnt = 0;
            last = true;
          }
        }
        const uint32_t last_mask = __ballot_sync(0xffffffffu, last);
        if (last_mask) __threadfence();
        if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(last_mask) : "memory");
      }
      __syncthreads();
      const uint32_t last_mask = lds_word(flag);
      stamp_item(2);
      if (last_mask != 0) {
        // Merge of the members this CTA completed: 32 rows x D/4 columns of float4 = 8 (row, column) items
        // per thread.  All of a thread's items advance together, two partials at a time, so 16 loads
        // are in flight per thread instead of one item's four (the merge of a whole sub-group took
        // up to 86 us as eight dependent passes per thread); every item still folds in chunk order.
        constexpr int IT = 32 * (D / 4) / 256;
        float Mx[IT], Ls[IT], a0[IT], a1[IT], a2[IT], a3[IT];
        long long pi0[IT];
        int na_i[IT];
        int na_max = 0;
#pragma unroll
        for (int u = 0; u < IT; ++u) {
          const int itx = tid + u * 256;
          const int r = itx / (D / 4);
          const int mm = r >> p.gp_shift, hd = r & gp_mask;
          const bool ok = mm < nmem && hd < g_valid && ((last_mask >> mm) & 1u);
          na_i[u] = ok ? member_chunks(mm) : 0;
          pi0[u] = ((long long)(b + (ok ? mm : 0)) * p.H + head0 + (ok ? hd : 0)) * p.max_splits;
          Mx[u] = -INFINITY;
          Ls[u] = a0[u] = a1[u] = a2[u] = a3[u] = 0.f;
          na_max = max(na_max, na_i[u]);
        }
        for (int s0 = 0; s0 < na_max; s0 += 2) {
          float2 ml[IT][2];
          float4 v[IT][2];
#pragma unroll
          for (int u = 0; u < IT; ++u) {
            const int dq = (tid + u * 256) % (D / 4);
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const int sidx = min(s0 + k, max(na_i[u] - 1, 0));
              ml[u][k] = __ldcg(p.part_ml + pi0[u] + sidx);
              v[u][k] = __ldcg(reinterpret_cast<const float4*>(p.part_o + (pi0[u] + sidx) * D + dq * 4));
            }
          }
#pragma unroll
          for (int u = 0; u < IT; ++u) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              if (s0 + k < na_i[u]) {
                const float Mn = fmaxf(Mx[u], ml[u][k].x);
                const float so = fast_exp2(Mx[u] - Mn), sn = fast_exp2(ml[u][k].x - Mn);
                Ls[u] = Ls[u] * so + ml[u][k].y * sn;
                a0[u] = a0[u] * so + v[u][k].x * sn;
                a1[u] = a1[u] * so + v[u][k].y * sn;
                a2[u] = a2[u] * so + v[u][k].z * sn;
                a3[u] = a3[u] * so + v[u][k].w * sn;
                Mx[u] = Mn;
              }
            }
          }
        }
#pragma unroll
        for (int u = 0; u < IT; ++u) {
          if (na_i[u] == 0) continue;
          const int itx = tid + u * 256;
          const int r = itx / (D / 4), dq = itx % (D / 4);
          const int mm = r >> p.gp_shift, hd = r & gp_mask;
          const float inv = 1.f / Ls[u];
          store_out4<T, D>(p, b + mm, head0 + hd, dq * 4, a0[u] * inv, a1[u] * inv, a2[u] * inv, a3[u] * inv);
        }
      }
      __syncthreads();  // the flag word is rewritten by the next item
      stamp_item(3);
      if (p.dbg_times != nullptr && tid == 0 && items_done < 4) p.dbg_times[192 + (blockIdx.x * 4 + items_done) * 5 + 4] = last_mask;
    }
    tc5::fence_before_sync();
  }

  __syncthreads();
  if (warp == 2) tc5::tmem_dealloc(tmem, C::TMEM_COLS);
}

template <typename T, int D>
int launch_group_tc5_d(const DecodeParams& p, const CUtensorMap& mq, long long batch, int block_size, int num_sms,
                       cudaStream_t s) {
  using C = Gt5Cfg<D>;
  auto kern = block_size == 16 ? paged_decode_group_tc5_kernel<T, D, true> : paged_decode_group_tc5_kernel<T, D, false>;
  static bool configured[64][2] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail_cuda((int)e, "cudaGetDevice");
  const int di = block_size == 16 ? 1 : 0;
  if (dev >= 64 || !configured[dev][di]) {
    if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM)) != cudaSuccess)
      return fail_cuda((int)e, "cudaFuncSetAttribute (group tcgen05)");
    if (dev < 64) configured[dev][di] = true;
  }
  // programmatic dependent of the pre-pass (its set-up overlaps that; it waits before reading the work list);
  // the split kernel follows as ITS programmatic dependent
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)num_sms);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = C::SMEM;
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  e = cudaLaunchKernelEx(&cfg, kern, mq, p, (int)batch);
  if (e != cudaSuccess) return fail_cuda((int)e, "paged_decode_group_tc5 launch");
  return 0;
}

template <typename T>
int launch_group_tc5(const DecodeParams& p, int head_dim, long long batch, int block_size, int num_sms, int dtype,
                     cudaStream_t s) {
  const int gp = 1 << p.gp_shift;
  CUtensorMap mq;
  const int r = make_group_q_map(&mq, p.q, dtype, head_dim, p.H, batch, gp, 32 / gp);
  if (r != 0) return fail_text("tensor map (q rows)", r);
  return head_dim == 64 ? launch_group_tc5_d<T, 64>(p, mq, batch, block_size, num_sms, s)
                        : launch_group_tc5_d<T, 128>(p, mq, batch, block_size, num_sms, s);
}

}  // namespace

int launch_decode_group_tc5(const DecodeParams& p, int head_dim, long long batch, int block_size, long long num_blocks,
                            int num_sms, int dtype, cudaStream_t s) {
  if (head_dim != 64 && head_dim != 128) return -1000;
  // dtype: MVL_DTYPE_F16 = 1, MVL_DTYPE_BF16 = 2
  (void)num_blocks;
  return dtype == 1 ? launch_group_tc5<__half>(p, head_dim, batch, block_size, num_sms, dtype, s)
                    : launch_group_tc5<__nv_bfloat16>(p, head_dim, batch, block_size, num_sms, dtype, s);
}

}  // namespace mvl
