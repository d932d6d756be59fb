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
// M = 128 accumulator tile, a 64-token K/V tile is the N (scores) and then the K (P.V) dimension:
//   S (128 x 64 fp32, TMEM) = Q (rows x D, smem)  . K_tile^T (64 tokens x D, smem)      D/16 tcgen05.mma
//   O (128 x D fp32, TMEM) += P (rows x 64, TMEM) . V_tile   (64 tokens x D, smem)      4 tcgen05.mma
// One thread issues all of it; rows 32..127 of the tile are padding the tensor pipe does not miss.
//
// Shape of the kernel: one CTA per SM, persistent over the work items (sub-group, kv head, shared chunk)
// that the device pre-pass listed (prefix_share.cu), 256 threads in roles that each walk the CTA's items on
// their own and meet only through mbarriers (no CTA-wide barrier between items):
//   warps 0, 4  softmax of the even / odd tiles, two independent streams (own maximum, sum, O accumulator):
//           thread r owns row r = TMEM lane r (member r >> log2 Gp, head r & (Gp - 1)); a score row arrives
//           with tcgen05.ld, P goes back over the first half of its S buffer as 16-bit pairs.  At the end of
//           an item warp 0 folds the two streams (it reads both accumulators) and writes the item's partials
//   warps 1, 3  K/V producers: a 6-stage ring of (K tile, V tile) = 6 x 32 KB filled with 16-byte cp.async
//           copies, rows located through the leader's block-table row (the members' rows agree on these
//           pages; entries prefetched in windows of 32), written in the 128-byte-swizzled layout the MMA
//           descriptors read; "full" barriers by cp.async.mbarrier.arrive, "empty" barriers armed by
//           tcgen05.commit.  They run into the next item as soon as the ring has room
//   warp 2  lane 0 issues every MMA from an event loop (Q.K up to three tiles ahead of P.V, three S buffers
//           in TMEM) and fetches the next item's Q rows by TMA into the other 4 KB slot of the Q tile
//   warps 5-7  finishers: count the members' arrivals for the item just written and merge the partials of
//           the members this CTA completed, while the next item streams; the CTA's last item is finished
//           by all 256 threads
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
  static constexpr int OFF_XML = OFF_BAR + 256;      // (m, l) of the odd softmax stream, two items deep
  static constexpr int SMEM = OFF_XML + 512 + 1024;  // + alignment slack
  // Three score buffers S0 S1 S2 of TPC columns, then O of the even tiles (D columns) and O of the odd tiles.
  // Three, so that Q.K of tile j+2 can be issued while tile j's softmax still runs: with two, a softmax warp
  // sat idle between its tiles for a P.V issue + Q.K round trip (0.5 us of every 1.5 us, measured) and the
  // K/V ring drained at one tile per 0.77 us, below the SM's share of HBM.
  static constexpr int SBUF = 3;
  static constexpr int O_COL = SBUF * TPC;
  static constexpr int TMEM_COLS = (SBUF * TPC + 2 * D) <= 256 ? 256 : 512;
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
  auto bar_q = [&](int slot) { return bar0 + 8 * slot; };                        // Q rows of an item landed (two slots)
  auto bar_kf = [&](int st) { return bar0 + 16 + 8 * st; };                     // K tile of a stage landed
  auto bar_vf = [&](int st) { return bar0 + 16 + 8 * (STAGES + st); };          // V tile landed
  auto bar_emp = [&](int st) { return bar0 + 16 + 8 * (2 * STAGES + st); };     // both consumed
  auto bar_sf = [&](int sb) { return bar0 + 16 + 8 * (3 * STAGES + sb); };      // S buffer holds a tile's scores
  auto bar_pr = [&](int sb) { return bar0 + 16 + 8 * (3 * STAGES + 3 + sb); };  // ... its P
  auto bar_o = [&](int sb) { return bar0 + 16 + 8 * (3 * STAGES + 6 + sb); };   // P.V into O[sb] retired
  const uint32_t bar_drain = bar0 + 16 + 8 * (3 * STAGES + 8);  // both O accumulators of an item read out
  const uint32_t bar_done = bar_drain + 8;                      // the item's partials are in global memory
  const uint32_t bar_fin = bar_drain + 16;                      // the finishers are through with the item
  const uint32_t tmem_slot = bar_drain + 24, flag = tmem_slot + 4;
  static_assert(16 + 8 * (3 * STAGES + 11) + 8 <= 256, "barrier block");
  const uint32_t s_xml = base + C::OFF_XML;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 64) tc5::prefetch_tensormap(&mq);
  if (warp == 2) {
    tc5::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc5::tmem_relinquish();
  }
  if (tid == 32) {
    mbar_init(bar_q(0), 1), mbar_init(bar_q(1), 1);
    mbar_init(bar_drain, 1), mbar_init(bar_done, 1), mbar_init(bar_fin, 1);
    for (int st = 0; st < STAGES; ++st) mbar_init(bar_kf(st), 64), mbar_init(bar_vf(st), 64), mbar_init(bar_emp(st), 1);
    for (int sb = 0; sb < C::SBUF; ++sb) mbar_init(bar_sf(sb), 1), mbar_init(bar_pr(sb), 1);
    mbar_init(bar_o(0), 1), mbar_init(bar_o(1), 1);
    mbar_fence_init();
  }
  // rows of the Q tile the loads never write: zeros (they only feed accumulator rows nobody reads)
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
  // ... and per CTA and item: start / partials written / counted / merged (called by one thread)
  auto stamp_item = [&](uint32_t n, int which) {
    if (p.dbg_times != nullptr && n < 4) {
      unsigned long long tnow;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tnow));
      p.dbg_times[192 + (blockIdx.x * 4 + n) * 5 + which] = tnow;
    }
  };
  const int gp = 1 << p.gp_shift, gp_mask = gp - 1;
  const int rows_box = 32;                          // (32 / gp) members x gp heads
  const int tiles = p.chunk / TPC;                  // a shared chunk is a whole chunk (multiple of 128): even, >= 2
  const uint32_t q_bytes = (uint32_t)C::PANELS * rows_box * 128;
  constexpr uint32_t HI = tc5::smem_desc_hi_sw128(1024);
  // An item's 32 Q rows take the first 4 KB of each panel of the Q tile; the next item's go behind them
  // (rows 32..63), so that they can be fetched while this item's Q.K still read theirs.  The 128-row
  // operand of an odd item starts at row 32 and runs into whatever follows the tile: rows nobody reads.
  constexpr uint32_t Q_SLOT_BYTES = 32 * 128;
  const int g_valid = p.G;

  // ---- which (sub-group, kv head, shared chunk) a work item is
  struct Item {
    int b, split, nmem, start, head0, x;
  };
  auto item_of = [&](long long item) -> Item {
    Item it;
    it.x = (int)(item % X);
    const int ci = (int)(item / X);
    int lo = 0, hi = wl_n;  // wl_cum[lo] <= ci < wl_cum[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (wl_cum[mid] <= ci) lo = mid; else hi = mid;
    }
    it.b = wl_b[lo];
    it.split = ci - wl_cum[lo];
    it.nmem = gsize[it.b];
    it.start = it.split * p.chunk;
    it.head0 = it.x * p.G;
    return it;
  };

  // ---- finishing an item: per member, the last of its chunk CTAs to arrive merges its partials (the split
  // kernel's protocol).  Run by a group of `nt` threads (thread `ft` of it) that meets on named barrier `bar_id`:
  // the three finisher warps while the item behind is streaming, the whole CTA for its last item.
  auto finish = [&](const Item& it, uint32_t n, int ft, int nt, int bar_id) {
    const int b = it.b, x = it.x, nmem = it.nmem, head0 = it.head0;
    auto group_sync = [&] { asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(nt) : "memory"); };
    if (p.counters != nullptr) {
      auto member_chunks = [&](int mm) -> int {
        long long c = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b + mm]
                                   : (long long)reinterpret_cast<const int*>(p.ctx)[b + mm];
        if (c > p.max_ctx) c = p.max_ctx;
        return (int)((c + p.chunk - 1) / p.chunk);
      };
      // lane m of the group's first warp counts member m's arrival (the members' atomics in parallel, not
      // one thread after the other: that loop cost 5 us per item)
      if (ft < 32) {
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
        const uint32_t lm = __ballot_sync(0xffffffffu, last);
        if (lm) __threadfence();
        if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(lm) : "memory");
      }
      group_sync();
      const uint32_t last_mask = lds_word(flag);
      if (ft == 0) stamp_item(n, 2);
      if (last_mask != 0) {
        // Merge of the members this CTA completed: 32 rows x D/4 columns of float4, IT (row, column) items
        // per thread and pass.  All of a thread's items advance together, two partials at a time, so 16
        // loads are in flight per thread instead of one item's four (the merge of a whole sub-group took
        // up to 86 us as eight dependent passes per thread); every item still folds in chunk order.
        constexpr int IT = 4, PB = 4, NI = 32 * (D / 4);
        for (int base_i = 0; base_i < NI; base_i += nt * IT) {
          float Mx[IT], Ls[IT], a0[IT], a1[IT], a2[IT], a3[IT];
          long long pi0[IT];
          int na_i[IT];
          int na_max = 0;
#pragma unroll
          for (int u = 0; u < IT; ++u) {
            const int itx = base_i + ft + u * nt;
            const int r = itx / (D / 4);
            const int mm = r >> p.gp_shift, hd = r & gp_mask;
            const bool ok = itx < NI && mm < nmem && hd < g_valid && ((last_mask >> mm) & 1u);
            na_i[u] = ok ? member_chunks(mm) : 0;
            pi0[u] = ((long long)(b + (ok ? mm : 0)) * p.H + head0 + (ok ? hd : 0)) * p.max_splits;
            Mx[u] = -INFINITY;
            Ls[u] = a0[u] = a1[u] = a2[u] = a3[u] = 0.f;
            na_max = max(na_max, na_i[u]);
          }
          for (int s0 = 0; s0 < na_max; s0 += PB) {
            float2 ml[IT][PB];
            float4 v[IT][PB];
#pragma unroll
            for (int u = 0; u < IT; ++u) {
              const int dq = (base_i + ft + u * nt) % (D / 4);
#pragma unroll
              for (int k = 0; k < PB; ++k) {
                const int sidx = min(s0 + k, max(na_i[u] - 1, 0));
                ml[u][k] = __ldcg(p.part_ml + pi0[u] + sidx);
                v[u][k] = __ldcg(reinterpret_cast<const float4*>(p.part_o + (pi0[u] + sidx) * D + dq * 4));
              }
            }
#pragma unroll
            for (int u = 0; u < IT; ++u) {
#pragma unroll
              for (int k = 0; k < PB; ++k) {
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
            const int itx = base_i + ft + u * nt;
            const int r = itx / (D / 4), dq = itx % (D / 4);
            const int mm = r >> p.gp_shift, hd = r & gp_mask;
            const float inv = 1.f / Ls[u];
            store_out4<T, D>(p, b + mm, head0 + hd, dq * 4, a0[u] * inv, a1[u] * inv, a2[u] * inv, a3[u] * inv);
          }
        }
      }
      group_sync();  // the flag word is rewritten by the next item
      if (ft == 0) {
        stamp_item(n, 3);
        if (p.dbg_times != nullptr && n < 4) p.dbg_times[192 + (blockIdx.x * 4 + n) * 5 + 4] = last_mask;
      }
    }
  };
  const uint32_t my_items = (uint32_t)((n_items - blockIdx.x + use_ctas - 1) / use_ctas);

  // Every role walks this CTA's items on its own, meeting the others only through mbarriers: the producers
  // are copying the next item's tiles while the last scores of this one are still being exponentiated, its
  // Q rows are already there, and counting / merging is the business of three otherwise idle warps.  (With
  // a __syncthreads() at the end of every item the ring ran dry, the partials were written, counted and
  // merged, and the ring filled again: 7.7 us of every 57.)  `gt` = tiles streamed before the item; all
  // roles count alike: ring stage g % STAGES, score buffer g % 3, and the phases that follow from them.
  if (warp == 1 || warp == 3) {
    // ================================================================ K/V producers (64 threads)
    // 16-byte cp.async copies straight into the 128-byte-swizzled tile layout the MMA descriptors
    // expect (row = token, 128 bytes per row and panel, chunk index XOR row % 8).  Measured first with
    // TMA boxes: with 16-slot pages a tile is 32 boxes of 2 KB, and the TMA unit's box rate capped the
    // kernel at 2.2 TB/s.  A thread owns chunk column `cc` of rows rg, rg + 4, ...; it arrives on the
    // stage's barriers with cp.async.mbarrier.arrive.noinc, i.e. when its copies have landed.
    const int pt = (warp == 1 ? 0 : 32) + lane;        // 0..63
    const int cc = pt % (C::PANELS * 8), rg = pt / (C::PANELS * 8);
    constexpr int RSTEP = 64 / (C::PANELS * 8);         // rows covered per pass of the 64 threads
    uint32_t gt = 0, n = 0;
    for (long long item = blockIdx.x; item < n_items; item += use_ctas, ++n) {
      const Item it = item_of(item);
      const int32_t* table = p.bt + (long long)it.b * p.bt_stride;
      const int bs_mask = (1 << p.bs_shift) - 1;
      const uint32_t row_pitch = (uint32_t)p.KVH * (D * 2);
      const char* kbase = reinterpret_cast<const char*>(p.kc) + (long long)it.x * (D * 2) + cc * 16;
      const char* vbase = reinterpret_cast<const char*>(p.vc) + (long long)it.x * (D * 2) + cc * 16;
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
      // The item's table entries travel in windows of 32, one window ahead of the copies: lane l holds
      // entry pgs + 32 w + l of the current window and of the next one, a row fetches its page with a
      // shuffle.  (One load per tile, issued when the tile came up, was measured as the ring's longest
      // link: every other tile missed L1 for its 32-byte table sector and the stage sat empty for ~1 us
      // before its first copy left.)  A tile's pages never straddle a window: TPC / block_size of them,
      // aligned to that count from the item's first page.
      const int pgs = it.start >> p.bs_shift;
      const int pg_last = (it.start + p.chunk - 1) >> p.bs_shift;
      auto load_window = [&](int w) -> uint32_t {
        const int e = pgs + w * 32 + lane;
        return e <= pg_last ? (uint32_t)__ldg(table + e) : 0u;
      };
      uint32_t e_cur = load_window(0), e_nxt = load_window(1);
      int win = 0;
      for (int j = 0; j < tiles; ++j) {
        const uint32_t g = gt + j;
        const int st = g % STAGES;
        const int tok0 = it.start + j * TPC;
        const int rel = (tok0 >> p.bs_shift) - pgs;
        if ((rel >> 5) != win) {  // at most one window further: a tile advances by TPC / block_size <= 4 pages
          e_cur = e_nxt;
          ++win;
          e_nxt = load_window(win + 1);
        }
        const int wi = rel & 31;
        if (g >= (uint32_t)STAGES) tc5::mbar_wait_bounded(bar_emp(st), ((g / STAGES) - 1) & 1u);
        if (pt == 0 && n == 0) stamp(j, 0);
        const uint32_t kdst = sK(st) + col_off, vdst = sV(st) + col_off;
        if (PAGE16) {
          // 16-slot pages (the reference's default, attention.py:483): page and slot-in-page of every row
          // this thread copies are compile-time functions of the unrolled row counter, so a copy costs an
          // add and the cp.async.  A producer warp has its scheduler to itself — nothing hides the
          // latency of a longer dependent address chain (measured: 23 instructions per copy, 4.4 us per tile).
          const char* kpage[TPC / 16];
#pragma unroll
          for (int pg = 0; pg < TPC / 16; ++pg)
            kpage[pg] = kbase + (unsigned long long)(__shfl_sync(0xffffffffu, e_cur, wi + pg) << 4) * row_pitch;
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
            const uint32_t e = __shfl_sync(0xffffffffu, e_cur, ((tok >> p.bs_shift) - pgs) & 31);
            const uint32_t slot = (e << p.bs_shift) + (uint32_t)(tok & bs_mask);
            const uint32_t off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + (uint32_t)(((cc & 7) ^ (r & 7)) << 4);
            cp_async16(kdst + off, row_ptr(kbase, slot, row_pitch), 16u);
          }
          asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar_kf(st)) : "memory");
#pragma unroll 8
          for (int r = rg; r < TPC; r += RSTEP) {
            const int tok = tok0 + r;
            const uint32_t e = __shfl_sync(0xffffffffu, e_cur, ((tok >> p.bs_shift) - pgs) & 31);
            const uint32_t slot = (e << p.bs_shift) + (uint32_t)(tok & bs_mask);
            const uint32_t off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + (uint32_t)(((cc & 7) ^ (r & 7)) << 4);
            cp_async16(vdst + off, row_ptr(vbase, slot, row_pitch), 16u);
          }
        }
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar_vf(st)) : "memory");
        if (pt == 0 && n == 0) stamp(j, 1);
      }
      gt += tiles;
    }
  } else if (warp == 2) {
    // ================================================================ MMA issuer (and Q loader)
    if (lane == 0) {
      auto load_q = [&](const Item& it, uint32_t n) {
        const uint32_t slot = n & 1u;
        mbar_expect_tx(bar_q(slot), q_bytes);
#pragma unroll
        for (int pn = 0; pn < C::PANELS; ++pn)
          tma_load_3d(sQ + pn * C::Q_PANEL_BYTES + slot * Q_SLOT_BYTES, &mq, pn * 64, it.head0, it.b, bar_q(slot));
      };
      // K/V tiles are written by cp.async (generic proxy) and read by the tensor cores (async proxy)
      auto proxy_fence = [] { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); };
      uint32_t gt = 0, n = 0;
      load_q(item_of(blockIdx.x), 0);
      for (long long item = blockIdx.x; item < n_items; item += use_ctas, ++n) {
        if (item + use_ctas < n_items) {
          // the other slot's last reader was item n - 1: its last Q.K has to have retired
          if (n > 0) tc5::mbar_wait_bounded(bar_sf((gt - 1) % C::SBUF), ((gt - 1) / C::SBUF) & 1u);
          load_q(item_of(item + use_ctas), n + 1);
        }
        tc5::mbar_wait_bounded(bar_q(n & 1u), (n >> 1) & 1u);
        const uint32_t q_lo = tc5::smem_desc_lo(sQ + (n & 1u) * Q_SLOT_BYTES, 16);
        auto issue_qk = [&](uint32_t g) {
          const uint32_t k_lo = tc5::smem_desc_lo(sK(g % STAGES), 16);
          const uint32_t s_t = tmem + (g % C::SBUF) * TPC;
#pragma unroll
          for (int ks = 0; ks < D / 16; ++ks) {
            const uint32_t qoff = ((ks >> 2) * C::Q_PANEL_BYTES + (ks & 3) * 32) >> 4;
            const uint32_t koff = ((ks >> 2) * C::PANEL_BYTES + (ks & 3) * 32) >> 4;
            tc5::mma_ss(s_t, tc5::desc64(q_lo + qoff, HI), tc5::desc64(k_lo + koff, HI), IDESC_QK, ks > 0);
          }
          tc5::commit(bar_sf(g % C::SBUF));
        };
        // Event loop: whichever of {K tile of the next Q.K landed, P and V of the next P.V ready} comes first
        // is issued first.  Q.K may run up to three tiles ahead of P.V (tile n's scores go where tile n - 3's
        // P was, and the tensor pipe keeps issue order).  The first P.V of an item overwrites an accumulator:
        // it also waits until the softmax warps have read the previous item's out of TMEM.
        int nq = 0, npv = 0;  // tiles whose Q.K / P.V have been issued
        bool o_free = n == 0;
        uint32_t idle = 0;
        while (npv < tiles) {
          bool did = false;
          if (npv < nq) {
            const uint32_t g = gt + npv;
            if (!o_free) o_free = tc5::mbar_test(bar_drain, (n - 1) & 1u);
            if (o_free && tc5::mbar_test(bar_pr(g % C::SBUF), (g / C::SBUF) & 1u) &&
                tc5::mbar_test(bar_vf(g % STAGES), (g / STAGES) & 1u)) {
              proxy_fence();
              tc5::fence_after_sync();
              if (n == 0) stamp(npv, 3);
              const uint32_t v_lo = tc5::smem_desc_lo(sV(g % STAGES), C::PANEL_BYTES);
              const uint32_t p_t = tmem + (g % C::SBUF) * TPC;
#pragma unroll
              for (int kk = 0; kk < TPC / 16; ++kk)
                tc5::mma_ts(tmem + C::O_COL + (npv & 1) * D, p_t + kk * 8, tc5::desc64(v_lo + kk * (2048 >> 4), HI),
                            IDESC_PV, (npv > 1 || kk > 0) ? 1u : 0u);
              tc5::commit(bar_emp(g % STAGES));
              tc5::commit(bar_o(npv & 1));
              ++npv;
              did = true;
            }
          }
          if (nq < tiles && nq < npv + C::SBUF) {
            const uint32_t g = gt + nq;
            if (tc5::mbar_test(bar_kf(g % STAGES), (g / STAGES) & 1u)) {
              proxy_fence();
              tc5::fence_after_sync();
              if (n == 0 && nq > 0) stamp(nq, 2);
              issue_qk(g);
              ++nq;
              did = true;
            }
          }
          if (did) idle = 0;
          else if (++idle > (1u << 28)) __trap();  // a lost completion is a programming error: do not hang the GPU
        }
        gt += tiles;
      }
    }
    __syncwarp();
  } else if (warp == 0 || warp == 4) {
    // ================================================================ softmax: two streams
    // Warp 0 takes the even tiles, warp 4 the odd ones (both reach TMEM lanes 0..31), each with its own
    // running maximum, sum and O accumulator: one warp doing every tile's 64 x 32 exponentials was the
    // slowest link of the tile pipeline (measured: 0.8-1.0 us of softmax + 0.25 us hand-off per 128
    // tokens against 1.2 us of HBM time).  The two streams are merged at the end of the item like two chunks.
    const int q = warp >> 2;
    const uint32_t lane_base = tmem;  // TMEM lanes 0..31
    const uint32_t o_addr = lane_base + C::O_COL + q * D;
    uint32_t gt = 0, n = 0;
    for (long long item = blockIdx.x; item < n_items; item += use_ctas, ++n) {
      const Item it = item_of(item);
      if (tid == 0) stamp_item(n, 0);
      float m_ref = 0.f, l = 0.f;
      for (int j = q; j < tiles; j += 2) {
        const uint32_t g = gt + j;
        tc5::mbar_wait_bounded(bar_sf(g % C::SBUF), (g / C::SBUF) & 1u);
        tc5::fence_after_sync();
        if (lane == 0 && n == 0) stamp(j, 4);
        group_softmax_step<T, D, TPC>(lane_base + (g % C::SBUF) * TPC, o_addr, j < 2, p.qscale, m_ref, l, bar_o(q),
                                      ((g >> 1) - 1) & 1u);
        tc5::fence_before_sync();
        __syncwarp();
        if (lane == 0 && n == 0) stamp(j, 5);
        if (lane == 0) tc5::mbar_arrive(bar_pr(g % C::SBUF));
      }
      // this stream's last product has retired (gt is even: a stream's tiles are every other g)
      const uint32_t g_last = gt + tiles - 2 + q;
      tc5::mbar_wait_bounded(bar_o(q), (g_last >> 1) & 1u);
      const uint32_t xml = s_xml + (n & 1u) * 256 + (uint32_t)lane * 8;  // two items deep: warp 4 runs ahead
      if (q == 1) {
        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(xml), "f"(m_ref * p.qscale), "f"(l) : "memory");
        tc5::fence_before_sync();
        asm volatile("bar.sync 1, 64;" ::: "memory");  // both softmax warps (64 threads): named barrier 1
      } else {
        tc5::mbar_wait_bounded(bar_o(1), ((g_last + 1) >> 1) & 1u);  // ... and the odd stream's
        asm volatile("bar.sync 1, 64;" ::: "memory");
        tc5::fence_after_sync();
        // Both accumulators sit in TMEM lanes 0..31, which this warp reaches: fold the odd stream's row into
        // the even one's like a second chunk and write the item's partial in the split kernel's layout.
        float mo, lo2;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(mo), "=f"(lo2) : "r"(xml) : "memory");
        const float me = m_ref * p.qscale;
        const float Mx = fmaxf(me, mo);
        const float se = fast_exp2(me - Mx), so = fast_exp2(mo - Mx);
        const float Lsum = l * se + lo2 * so;
        const int m = lane >> p.gp_shift, hd = lane & gp_mask;
        const bool row_ok = m < it.nmem && hd < g_valid;
        const long long pi = ((long long)(it.b + m) * p.H + it.head0 + hd) * p.max_splits + it.split;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t oe[32], oo[32];
          tc5::tmem_ld32(o_addr + c * 32, oe);
          tc5::tmem_ld32(o_addr + D + c * 32, oo);
          tc5::tmem_wait_ld();
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<float4*>(p.part_o + pi * D + c * 32 + i) =
                  make_float4(__uint_as_float(oe[i]) * se + __uint_as_float(oo[i]) * so,
                              __uint_as_float(oe[i + 1]) * se + __uint_as_float(oo[i + 1]) * so,
                              __uint_as_float(oe[i + 2]) * se + __uint_as_float(oo[i + 2]) * so,
                              __uint_as_float(oe[i + 3]) * se + __uint_as_float(oo[i + 3]) * so);
          }
        }
        if (row_ok) p.part_ml[pi] = make_float2(Mx, Lsum);
        tc5::fence_before_sync();
        __threadfence();
        __syncwarp();
        if (lane == 0) {
          tc5::mbar_arrive(bar_drain);  // the next item's first products may overwrite the accumulators
          stamp_item(n, 1);
          if (n + 1 < my_items) {  // hand the item to the finishers (its last one is finished by the whole CTA)
            if (n > 0) tc5::mbar_wait_bounded(bar_fin, (n - 1) & 1u);  // they are through with the one before
            tc5::mbar_arrive(bar_done);
          }
        }
        __syncwarp();
      }
      gt += tiles;
    }
  } else if (warp >= 5) {
    // ================================================================ finishers (96 threads): all items but the last
    const int ft = tid - 160;
    uint32_t n = 0;
    for (long long item = blockIdx.x; n + 1 < my_items; item += use_ctas, ++n) {
      const Item it = item_of(item);
      tc5::mbar_wait_bounded(bar_done, n & 1u);
      finish(it, n, ft, 96, 2);
      if (ft == 0) tc5::mbar_arrive(bar_fin);
    }
  }

  // ---- the CTA's last item is finished by everybody: nothing is left to overlap it with
  tc5::fence_before_sync();
  __syncthreads();
  finish(item_of(blockIdx.x + (long long)(my_items - 1) * use_ctas), my_items - 1, tid, 256, 3);
  tc5::fence_before_sync();
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
