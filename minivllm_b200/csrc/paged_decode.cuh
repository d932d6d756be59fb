// paged_decode.cuh — split-KV paged GQA decode attention for sm_100a (kernel templates).
//
// Replaces the reference's Triton kernel paged_attention_decode_kernel
// (reference src/myvllm/layers/attention.py:283-415): same math, different machine mapping.
//
// Work decomposition
//   grid = (num_kv_heads * head_tiles, batch, max_splits).  One CTA owns the tokens
//   [split*chunk, min(ctx, (split+1)*chunk)) of one sequence for ONE kv head and ALL query heads of
//   its GQA group (GT heads per tile), so every K/V byte is fetched from HBM once per group instead
//   of once per query head (the reference's grid (batch, num_heads) re-reads it G times,
//   attention.py:455).  CTAs whose chunk starts past the sequence's length exit at once, which is
//   how ragged batches are handled under a host-fixed (graph-capturable) grid.
//
// Inside a CTA
//   NW warps, each streaming its own interleaved tiles of TR rows (rows = tokens) through a private
//   3-slot cp.async ring in shared memory (K tile, V tile, next K tile ... each 4 KB, 16-byte
//   cp.async.cg, coalesced 128-bit lanes).  No block-level barrier in the streaming loop.
//   A lane holds one 16-byte chunk of a row (VEC elements): LPR lanes cover a row, a warp covers
//   RPP rows per pass and R passes per tile.  q (pre-scaled by scale*log2e) and the output
//   accumulators live in registers as packed fp32 pairs; all FMAs are FFMA2.
//   Scores: per-lane partial dot products are transposed through a swizzled shared buffer and summed
//   so each lane owns VPL complete scores; the running maximum is only refreshed when a score exceeds
//   it by more than 8 (log2 units), so the common path has no cross-lane max reduction.
//   Probabilities are published to shared memory as (p,p) pairs and consumed by the P.V pass.
//   At the end the NW warps' (m, l, o) states are merged through shared memory and the CTA writes
//   either the final row (sequence fits one chunk) or an fp32 partial for the combine kernel.
#pragma once
#include <cuda.h>  // CUtensorMap (type only; the encoder is fetched from the driver at run time)

#include "common.cuh"

namespace mvl {

struct DecodeParams {
  // Result destinations: outs[0..n_out) are (batch, H_out, D) tensors; this launch's heads land at
  // head offset head_off in each.  n_out = 1, H_out = H, head_off = 0 is the plain single-GPU case;
  // with head sharding outs[] are the peer-mapped full-head buffers of all ranks, so the result is
  // all-gathered by the stores themselves (NVLink P2P) instead of by a separate collective.
  void* outs[8];
  int n_out, H_out, head_off;
  // Optional completion signal for the peer-store gather (peer_sync.cu): once the launch's last CTA
  // has stored its rows, this rank's step count is bumped and written to slot [my_rank] of every
  // rank's control block (system-scope release), where mvl_peer_wait polls it.  Control block =
  // 16 x u32: [0..7] step counts announced by ranks 0..7, [8] this rank's own count, [9] CTA
  // arrival counter, [10] wait time-out marker.
  // 1: the launch carries a programmatic dependency on the kernel in front of it (MVL_DECODE_OVERLAP_PREV):
  // context lengths, block tables and K/V rows older than each sequence's newest token may be read
  // before griddepcontrol.wait.  0: ordinary stream order.
  int overlap_prev;
  // 1: the tcgen05 group kernel (prefix sharing, wide form) was launched in FRONT of this split kernel and
  // this launch is its programmatic dependent: the two touch disjoint chunks, so the split kernel's CTAs
  // start wherever a group CTA has left an SM, do not wait at entry, and wait for that grid's completion
  // only before they exit (so that this kernel's completion implies that one's).
  int beside_group;
  // Diagnostics (mvl_debug_set_times): per CTA of the stream kernel four %globaltimer stamps —
  // kernel entry, piece known, first tile landed, exit — and the SM it ran on.  nullptr in normal use.
  unsigned long long* dbg_times;
  unsigned* sync_local;     // this rank's control block; nullptr = no signalling
  unsigned* sync_peers[8];  // every rank's control block as mapped here (entry my_rank = sync_local)
  int my_rank;
  // Optional fusion of the step's KV-cache write (store_kvcache of the sequence's newest token):
  // the CTA whose chunk holds token ctx-1 writes that row for its kv head before it streams it.
  const void* k_new;        // (batch, KVH, D), token strides in bytes; nullptr = no fused store
  const void* v_new;
  const void* slot_map;     // (batch,) int32/int64, slot of token ctx-1 (or -1)
  long long k_new_stride, v_new_stride, num_slots;
  int slot_is_i64;
  // Optional fusion of the combine: per (sequence, head tile) arrival counters (zero before the
  // call, left zero after it); the last chunk's CTA to arrive merges the partials itself.
  int* counters;            // nullptr = separate combine kernel
  // Optional prefix sharing (tensor-core kernel, G < 8): sequences whose block tables alias are
  // grouped by a device pre-pass (prefix_share.cu); for chunks every member of a sub-group shares,
  // the leader's CTA scores all members against ONE copy of the K/V tiles (the instruction's 8
  // columns hold 8/Gp sequences x Gp heads).  meta = leader[B], gsize[B], gshare[B], lcpt[B].
  const int* share_meta;    // nullptr = every sequence on its own
  int gp_shift;             // log2(Gp), Gp = heads per sequence in the 8 columns (8 when not sharing)
  int mm_narrow;            // sequences the split kernel packs into its 8 columns = 8 / Gp
  int wide;                 // 1: sub-groups hold up to 4 * mm_narrow sequences and the chunks shared by
                            // more than mm_narrow of them are left to paged_decode_group_mma_kernel
  const void* q;
  const void* kc;
  const void* vc;
  const int32_t* bt;
  const void* ctx;
  float* part_o;   // [batch][H][max_splits][D] fp32, un-normalised
  float2* part_ml; // [batch][H][max_splits] (m in log2 units, l)
  long long bt_stride;
  long long max_ctx;  // max_num_blocks * block_size
  int ctx_is_i64;
  int H, KVH, G;
  int n_gt;       // head tiles per kv head = ceil(G / GT)
  int bs_shift;   // log2(block_size)
  int chunk;      // tokens per split
  int max_splits;
  float qscale;      // scale * log2(e)
};

// Store 4 consecutive result elements of (sequence b, local head) to every destination.
template <typename T, int D>
__device__ __forceinline__ void store_out4(const DecodeParams& p, long long b, int head, int d0,
                                           float a, float b1, float c, float d) {
  const long long off = (b * p.H_out + p.head_off + head) * D + d0;
#pragma unroll 1
  for (int r = 0; r < p.n_out; ++r) store4<T>(reinterpret_cast<T*>(p.outs[r]) + off, a, b1, c, d);
}
template <typename T, int D>
__device__ __forceinline__ void store_out2(const DecodeParams& p, long long b, int head, int d0,
                                           float a, float b1) {
  const long long off = (b * p.H_out + p.head_off + head) * D + d0;
#pragma unroll 1
  for (int r = 0; r < p.n_out; ++r) store2<T>(reinterpret_cast<T*>(p.outs[r]) + off, a, b1);
}

template <typename T, int D_, int GT_>
struct DecodeCfg {
  static constexpr int D = D_;
  static constexpr int GT = GT_;
  static constexpr int NW = 4;                  // warps per CTA
  static constexpr int ES = sizeof(T);
  static constexpr int VEC = Chunk<T>::VEC;     // elements per 16-byte chunk
  static constexpr int NP = Chunk<T>::NP;       // fp32 pairs per chunk
  static constexpr int LPR = D / VEC;           // lanes per row
  static constexpr int RPP = 32 / LPR;          // rows per pass
  static constexpr int R = 8;                   // passes per tile
  static constexpr int TR = R * RPP;            // rows per warp tile
  static constexpr int ROWB = D * ES;           // bytes per row of one kv head
  static constexpr int TILEB = TR * ROWB;       // bytes per K (or V) tile = R * 512
  // Score phase granularity: the GT heads are scored GK at a time (NSUB sub-phases, q from shared
  // memory when NSUB > 1); the P.V phase always carries all GT heads.  GK = 4 for 8-head tiles fits
  // 3 CTAs per SM (167 registers) but measured SLOWER on B200 than scoring all 8 heads at once with
  // 2 CTAs per SM (C4: 6.54 ms vs 6.12 ms): the re-read and re-conversion of K and the second
  // transpose cost more issue slots than the extra warps recover.  So GK = GT.
  // (The one exception: 16-bit head_dim 64 would own 8 scores per lane; it uses two sub-phases.)
  static constexpr int GK = (GT_ >= 8 && LPR == 8) ? 4 : GT_;
  static constexpr int NSUB = GT_ / GK;
  static constexpr int NV = R * GK;             // partial scores per lane, tile and sub-phase
  static constexpr int NQ = NV / 4;             // ... in 16-byte quads
  static constexpr int VPL = (NV / LPR) > 0 ? (NV / LPR) : 1;  // complete scores per lane and sub-phase
  static constexpr int NSLOT = VPL * NSUB;      // complete scores per lane and tile
  static constexpr int XB_BYTES = 32 * NV * 4;  // transpose buffer
  static constexpr int PB_BYTES = ((TR * GT * 8 + 255) / 256) * 256;  // (p,p) pairs, 256-B granular
  static constexpr int RING = 3;                // ring slots; RING-1 half tiles in flight
  static constexpr int BAR_BYTES = 256;         // RING mbarriers (TMA variant), 256-B granular
  static constexpr int WARP_SMEM = RING * TILEB + XB_BYTES + PB_BYTES + BAR_BYTES;
  static constexpr int QB_BYTES = (NSUB > 1) ? GT * D * 4 : 0;  // CTA-wide pre-scaled q (fp32)
  static constexpr int SMEM = NW * WARP_SMEM + QB_BYTES;
  static constexpr int MRG_BYTES = RPP * GT * D * 4 + GT * 8;  // merge area, aliases the ring
  static_assert(LPR >= 8 && LPR <= 32, "head_dim / dtype combination not supported");
  static_assert(MRG_BYTES <= RING * TILEB, "merge area must fit in the ring");
  static_assert(NV % 4 == 0 && NQ >= 2, "NV must be a multiple of 8");
  static_assert(VPL == 1 || VPL == 2 || VPL == 4, "unsupported scores-per-lane");
  static_assert(VPL <= GK, "a lane's scores of one sub-phase must belong to one row");
  // the XOR swizzle of the transpose buffer needs rows aligned to their size (<= 256 B): every
  // region size is a multiple of 256 and the kernel uses no static shared memory
  static_assert((RING * TILEB) % 256 == 0 && XB_BYTES % 256 == 0 && WARP_SMEM % 256 == 0, "alignment");
};

// XOR swizzle of the transpose buffer: lane L stores quad k at quad (k ^ key(L)).
template <int NQ>
__device__ __forceinline__ int xb_key(int L) {
  if (NQ >= 8) return L & 7;
  if (NQ == 4) return (L >> 1) & 3;
  return (L >> 2) & 1;  // NQ == 2
}

// UNIFORM: every warp tile lies inside one page (block_size % TR == 0), so one block-table entry
// addresses the whole tile; otherwise each row looks up its own page.
// TMA: K/V tiles are fetched with ONE 3-D tiled TMA load per tile (cp.async.bulk.tensor over the
// tensor map [num_slots][KVH][D], box [TR][1][D]) issued by lane 0 and completed on a per-slot
// mbarrier, instead of 8 x 32 lane-wide cp.async (needs UNIFORM).  Same ring, same math.
template <typename T, int D, int GT, bool UNIFORM, bool TMA>
__global__ void __launch_bounds__(DecodeCfg<T, D, GT>::NW * 32, (DecodeCfg<T, D, GT>::GK >= 8 ? 2 : 3))
paged_decode_split_kernel(const DecodeParams p, const __grid_constant__ CUtensorMap tm_k,
                          const __grid_constant__ CUtensorMap tm_v) {
  static_assert(!TMA || UNIFORM, "the TMA variant needs whole tiles inside one page");
  using C = DecodeCfg<T, D, GT>;
  constexpr int NP = C::NP, R = C::R, RPP = C::RPP, LPR = C::LPR, TR = C::TR, NV = C::NV,
                NQ = C::NQ, VPL = C::VPL, NW = C::NW, ROWB = C::ROWB, TILEB = C::TILEB,
                RING = C::RING, GK = C::GK, NSUB = C::NSUB, NSLOT = C::NSLOT;
  extern __shared__ __align__(1024) unsigned char smem[];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = lane / LPR, c = lane % LPR;
  const int kvh = blockIdx.x / p.n_gt, gt = blockIdx.x % p.n_gt;
  // y = sequence, z = chunk: CTAs of the same chunk index of all sequences are scheduled together,
  // so sequences whose block tables alias (shared prefixes) read the same pages at about the same
  // time and all but the first reader hit L2.
  const int b = blockIdx.y, split = blockIdx.z;
  const int head0 = kvh * p.G + gt * GT;
  const int g_valid = min(GT, p.G - gt * GT);

  long long ctx_ll = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b]
                                  : (long long)reinterpret_cast<const int*>(p.ctx)[b];
  if (ctx_ll > p.max_ctx) ctx_ll = p.max_ctx;
  const int ctx = (int)ctx_ll;

  if (ctx <= 0) {  // graph-padding row: defined as zeros (reference gives 0/0)
    if (split == 0)
      for (int i = threadIdx.x; i < g_valid * (D / 4); i += NW * 32) {
        const int g = i / (D / 4), dq = i % (D / 4);
        store_out4<T, D>(p, b, head0 + g, dq * 4, 0.f, 0.f, 0.f, 0.f);
      }
    return;
  }
  const int start = split * p.chunk;
  if (start >= ctx) return;
  const int end = min(ctx, start + p.chunk);
  const int n_act = (ctx + p.chunk - 1) / p.chunk;

  if (p.k_new != nullptr && ctx - 1 >= start && ctx - 1 < end) {  // CTA-uniform
    const long long slot = p.slot_is_i64 ? reinterpret_cast<const long long*>(p.slot_map)[b]
                                         : (long long)reinterpret_cast<const int*>(p.slot_map)[b];
    if (slot >= 0 && slot < p.num_slots) {
      constexpr int CPR = C::ROWB / 16;  // 16-byte chunks per row of one kv head
      if (threadIdx.x < 2 * CPR) {
        const int isv = threadIdx.x / CPR, ch = threadIdx.x % CPR;
        const char* src = reinterpret_cast<const char*>(isv ? p.v_new : p.k_new) +
                          b * (isv ? p.v_new_stride : p.k_new_stride) + (long long)kvh * C::ROWB + ch * 16;
        char* dst = reinterpret_cast<char*>(const_cast<void*>(isv ? p.vc : p.kc)) +
                    (slot * p.KVH + kvh) * C::ROWB + ch * 16;
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
      }
      __threadfence();  // the row is read back below through cp.async (L2)
    }
    __syncthreads();
  }

  const uint32_t wbase = smem_u32(smem) + warp * C::WARP_SMEM;
  const uint32_t ring = wbase;
  const uint32_t xbuf = wbase + C::RING * TILEB;
  const uint32_t pbuf = xbuf + C::XB_BYTES;
  const uint32_t bars = pbuf + C::PB_BYTES;  // RING mbarriers of this warp (TMA variant)
  if (TMA) {
    if (lane == 0) {
#pragma unroll
      for (int s = 0; s < RING; ++s) mbar_init(bars + s * 8, 1);
      mbar_fence_init();
    }
    __syncwarp();
  }

  // ---- q for the GT heads of this tile, as packed fp32 pairs scaled by scale*log2(e):
  // in registers when all heads are scored at once, else in shared memory (qbuf, [head][D] fp32)
  u64 q2[GK][NP];
  const uint32_t qbuf = smem_u32(smem) + NW * C::WARP_SMEM;
  {
    const u64 qs = pack2(p.qscale, p.qscale);
    if (NSUB == 1) {
#pragma unroll
      for (int g = 0; g < GK; ++g) {
        uint4 raw = make_uint4(0, 0, 0, 0);
        if (g < g_valid)
          raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.q) +
                                                ((long long)b * p.H + head0 + g) * D + c * C::VEC);
        Chunk<T>::widen(raw, q2[g]);
#pragma unroll
        for (int e = 0; e < NP; ++e) q2[g][e] = fmul2(q2[g][e], qs);
      }
    } else {
      for (int i = threadIdx.x; i < GT * LPR; i += NW * 32) {  // one 16-byte chunk of q each
        const int g = i / LPR, cc = i % LPR;
        uint4 raw = make_uint4(0, 0, 0, 0);
        if (g < g_valid)
          raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.q) +
                                                ((long long)b * p.H + head0 + g) * D + cc * C::VEC);
        u64 w2[NP];
        Chunk<T>::widen(raw, w2);
#pragma unroll
        for (int e = 0; e < NP; e += 2) {
          float a0, a1, a2, a3;
          unpack2(fmul2(w2[e], qs), a0, a1);
          unpack2(fmul2(w2[e + 1], qs), a2, a3);
          sts128(qbuf + (g * D + cc * C::VEC + e * 2) * 4,
                 make_uint4(__float_as_uint(a0), __float_as_uint(a1), __float_as_uint(a2),
                            __float_as_uint(a3)));
        }
      }
      __syncthreads();
    }
  }

  // ---- per-lane ownership of complete scores: slot s = sub*VPL + i <-> (pass j, head g), row rr
  int rr_s[NSLOT], g_s[NSLOT];
  bool ok_s[NSLOT];
#pragma unroll
  for (int s = 0; s < NSLOT; ++s) {
    const int idx = c * VPL + (s % VPL);
    ok_s[s] = idx < NV;
    g_s[s] = (s / VPL) * GK + idx % GK;
    rr_s[s] = (idx / GK) * RPP + h;
  }
  const int idx0 = (c * VPL < NV) ? c * VPL : 0;  // first owned score (clamped for idle lanes)

  // ---- this warp's tiles: tile = warp, warp + NW, ...
  const int ntiles = (end - start + TR - 1) / TR;
  const int my_nt = (warp < ntiles) ? (ntiles - warp + NW - 1) / NW : 0;
  const uint32_t row_pitch = (uint32_t)p.KVH * ROWB;  // bytes between consecutive slots
  const int bs_mask = (1 << p.bs_shift) - 1;
  const int32_t* bt_row = p.bt + (long long)b * p.bt_stride;
  const char* kbase = reinterpret_cast<const char*>(p.kc) + (long long)kvh * ROWB + c * 16;
  const char* vbase = reinterpret_cast<const char*>(p.vc) + (long long)kvh * ROWB + c * 16;
  const uint32_t lane_off = h * ROWB + c * 16;  // this lane's byte inside a tile pass

  // Block-table entries are fetched one tile ahead of the cp.async that needs them, so the
  // table load's latency never sits in front of an issue.  blk_a: entry of the tile whose half
  // tiles are issued next; blk_b: entry of the tile after it.
  auto table_entry = [&](int i) -> uint32_t {
    return (UNIFORM && i < my_nt)
               ? (uint32_t)__ldg(bt_row + ((start + (warp + i * NW) * TR) >> p.bs_shift))
               : 0u;
  };
  uint32_t blk_a = table_entry(0), blk_b = table_entry(1);

  // Issue the n-th half tile (even: K of tile n/2, odd: V) into ring slot `slot`; always commits.
  auto issue = [&](int n, int slot) {
    if (n < 2 * my_nt) {
      const int tstart = start + (warp + (n >> 1) * NW) * TR;
      const char* base = (n & 1) ? vbase : kbase;
      const uint32_t dst = ring + slot * TILEB + lane_off;
      if (TMA) {
        if (lane == 0) {
          const uint32_t bar = bars + slot * 8;
          mbar_expect_tx(bar, TILEB);
          tma_load_3d(ring + slot * TILEB, (n & 1) ? &tm_v : &tm_k, 0, kvh,
                      (int)((blk_a << p.bs_shift) + (tstart & bs_mask)), bar);
        }
      } else if (UNIFORM) {
        const uint32_t row0 = (blk_a << p.bs_shift) + (tstart & bs_mask) + h;
        if (tstart + TR <= end) {  // whole tile valid: no per-row predicate
#pragma unroll
          for (int j = 0; j < R; ++j)
            cp_async16(dst + j * RPP * ROWB, row_ptr(base, row0 + j * RPP, row_pitch), 16u);
        } else {
#pragma unroll
          for (int j = 0; j < R; ++j)
            cp_async16(dst + j * RPP * ROWB, row_ptr(base, row0 + j * RPP, row_pitch),
                       (tstart + j * RPP + h < end) ? 16u : 0u);
        }
      } else {
#pragma unroll
        for (int j = 0; j < R; ++j) {
          const int t = tstart + j * RPP + h;
          const bool valid = t < end;
          const int tt = valid ? t : end - 1;
          const uint32_t blk = (uint32_t)__ldg(bt_row + (tt >> p.bs_shift));
          cp_async16(dst + j * RPP * ROWB,
                     row_ptr(base, (blk << p.bs_shift) + (tt & bs_mask), row_pitch), valid ? 16u : 0u);
        }
      }
    }
    if (!TMA) cp_async_commit();
    if (UNIFORM && (n & 1)) {  // both halves of tile n/2 issued: advance the table prefetch
      blk_a = blk_b;
      blk_b = table_entry((n >> 1) + 2);
    }
  };

#pragma unroll 1
  for (int n = 0; n < RING - 1; ++n) issue(n, n);

  float m_all[GT];
  u64 o2[GT][NP];
  float m_slot[NSLOT], l_slot[NSLOT];
#pragma unroll
  for (int g = 0; g < GT; ++g) {
    m_all[g] = -INFINITY;
#pragma unroll
    for (int e = 0; e < NP; ++e) o2[g][e] = 0ull;
  }
#pragma unroll
  for (int i = 0; i < NSLOT; ++i) {
    m_slot[i] = -INFINITY;
    l_slot[i] = 0.f;
  }

  // loop-invariant shared-memory addresses, kept to a handful of registers
  const uint32_t xw = xbuf + lane * (NV * 4) + (xb_key<NQ>(lane) << 4);     // writer: xw ^ (k << 4)
  const uint32_t xr = xbuf + h * (LPR * NV * 4) + ((idx0 >> 2) << 4) + (idx0 & 3) * 4;  // reader
  // (p,p) pairs live as [row group h][pass j][head g]; a lane's VPL scores of one sub-phase are
  // consecutive heads of one pass, so they are stored with one or two 16-byte stores
  const uint32_t pw = pbuf + (h * R * GT + (idx0 / GK) * GT + idx0 % GK) * 8;
  const uint32_t pr = pbuf + h * R * GT * 8;                                // pairs of my row group

  int slot_c = 0;  // ring slot of the half tile being consumed
  uint32_t phase = 0;  // mbarrier parity of that slot's current use (TMA variant)
#pragma unroll 1
  for (int n = 0; n < 2 * my_nt; ++n) {
    if (TMA) {
      __syncwarp();  // every lane is done with the slot about to be refilled
      issue(n + RING - 1, slot_c == 0 ? RING - 1 : slot_c - 1);
      mbar_wait(bars + slot_c * 8, phase);
      const int ts = start + (warp + (n >> 1) * NW) * TR;
      if (ts + TR > end) {  // partial last tile: rows past the end hold whatever the page holds
#pragma unroll
        for (int j = 0; j < R; ++j)
          if (ts + j * RPP + h >= end)
            sts128(ring + slot_c * TILEB + lane_off + j * RPP * ROWB, make_uint4(0, 0, 0, 0));
        fence_proxy_async();  // before a later TMA load refills these bytes
      }
    } else {
      cp_async_wait<RING - 2>();
      __syncwarp();
      issue(n + RING - 1, slot_c == 0 ? RING - 1 : slot_c - 1);
    }
    const uint32_t tile = ring + slot_c * TILEB + lane_off;
    slot_c = (slot_c + 1 == RING) ? 0 : slot_c + 1;
    if (TMA && slot_c == 0) phase ^= 1;
    if (!(n & 1)) {
      // ================================================================ K: scores
      const int tstart = start + (warp + (n >> 1) * NW) * TR;
      float sc[NSLOT];
#pragma unroll
      for (int sub = 0; sub < NSUB; ++sub) {
        if (NSUB > 1) {
          if (sub > 0) __syncwarp();  // the previous sub-phase's gather has read the transpose buffer
#pragma unroll
          for (int g = 0; g < GK; ++g)
#pragma unroll
            for (int e = 0; e < NP; e += 2) {
              uint4 v = lds128(qbuf + ((sub * GK + g) * D + c * C::VEC + e * 2) * 4);
              q2[g][e] = pack2(__uint_as_float(v.x), __uint_as_float(v.y));
              q2[g][e + 1] = pack2(__uint_as_float(v.z), __uint_as_float(v.w));
            }
        }
        {
          float sv[NV];
#pragma unroll
          for (int j = 0; j < R; ++j) {
            u64 k2[NP];
            Chunk<T>::widen(lds128(tile + j * RPP * ROWB), k2);
#pragma unroll
            for (int g = 0; g < GK; ++g) {
              u64 acc = 0ull;
#pragma unroll
              for (int e = 0; e < NP; ++e) acc = ffma2(q2[g][e], k2[e], acc);
              sv[j * GK + g] = hsum2(acc);
            }
            // publish the quads of partial scores completed by this pass (keeps few of them live)
#pragma unroll
            for (int k = (j * GK) / 4; k < ((j + 1) * GK) / 4; ++k)
              sts128(xor_imm(xw, k << 4),
                     make_uint4(__float_as_uint(sv[4 * k]), __float_as_uint(sv[4 * k + 1]),
                                __float_as_uint(sv[4 * k + 2]), __float_as_uint(sv[4 * k + 3])));
          }
        }
        __syncwarp();
        // gather: lane (h,c) sums the LPR partials of its VPL scores (source lane h*LPR+cc, whose
        // swizzle key depends only on cc)
        {
          constexpr int NPAIR = (VPL + 1) / 2;
          u64 accp[NPAIR];
          float acc1 = 0.f;
#pragma unroll
          for (int i = 0; i < NPAIR; ++i) accp[i] = 0ull;
          constexpr int NK = (NQ >= 8) ? 8 : NQ;       // distinct swizzle keys
          constexpr int KSH = (NQ >= 8) ? 0 : (NQ == 4 ? 1 : 2);  // key(cc) = (cc >> KSH) & (NK-1)
#pragma unroll
          for (int kk = 0; kk < NK; ++kk) {
            const uint32_t a = xor_imm(xr, kk << 4);
#pragma unroll
            for (int cc = 0; cc < LPR; ++cc) {
              if (((cc >> KSH) & (NK - 1)) != kk) continue;
              const uint32_t at = a + cc * (NV * 4);
              if (VPL == 1) {
                acc1 += __uint_as_float(lds32(at));
              } else if (VPL == 2) {
                uint2 v = lds64(at);
                accp[0] = fadd2(accp[0], pack2(__uint_as_float(v.x), __uint_as_float(v.y)));
              } else {
                uint4 v = lds128(at);
                accp[0] = fadd2(accp[0], pack2(__uint_as_float(v.x), __uint_as_float(v.y)));
                accp[1 % NPAIR] = fadd2(accp[1 % NPAIR], pack2(__uint_as_float(v.z), __uint_as_float(v.w)));
              }
            }
          }
          if (VPL == 1) {
            sc[sub * VPL] = acc1;
          } else {
#pragma unroll
            for (int i = 0; i < NPAIR; ++i)
              unpack2(accp[i], sc[sub * VPL + (2 * i) % VPL], sc[sub * VPL + (2 * i + 1) % VPL]);
          }
        }
      }
      // ---- online softmax with a lazily refreshed maximum
      bool need = false;
#pragma unroll
      for (int i = 0; i < NSLOT; ++i) {
        const bool valid = ok_s[i] && (tstart + rr_s[i] < end);
        sc[i] = valid ? sc[i] : -INFINITY;
        need |= sc[i] > m_slot[i] + 8.f;
      }
      if (__any_sync(0xffffffffu, need)) {
        // rare: publish raw scores as [row][head], every lane takes the per-head maxima
#pragma unroll
        for (int i = 0; i < NSLOT; ++i)
          if (ok_s[i]) sts32(pbuf + (rr_s[i] * GT + g_s[i]) * 4, __float_as_uint(sc[i]));
        __syncwarp();
        float mx[GT];
#pragma unroll
        for (int g = 0; g < GT; ++g) mx[g] = m_all[g];
#pragma unroll
        for (int qd = 0; qd < TR * GT / 4; ++qd) {
          uint4 v = lds128(pbuf + qd * 16);
          mx[(4 * qd + 0) % GT] = fmaxf(mx[(4 * qd + 0) % GT], __uint_as_float(v.x));
          mx[(4 * qd + 1) % GT] = fmaxf(mx[(4 * qd + 1) % GT], __uint_as_float(v.y));
          mx[(4 * qd + 2) % GT] = fmaxf(mx[(4 * qd + 2) % GT], __uint_as_float(v.z));
          mx[(4 * qd + 3) % GT] = fmaxf(mx[(4 * qd + 3) % GT], __uint_as_float(v.w));
        }
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          const float alpha = fast_exp2(m_all[g] - mx[g]);  // m_all = -inf -> 0
          const u64 a2 = pack2(alpha, alpha);
#pragma unroll
          for (int e = 0; e < NP; ++e) o2[g][e] = fmul2(o2[g][e], a2);
#pragma unroll
          for (int i = 0; i < NSLOT; ++i)
            if (g_s[i] == g) {
              l_slot[i] *= alpha;
              m_slot[i] = mx[g];
            }
          m_all[g] = mx[g];
        }
        __syncwarp();
      }
      float pr_v[NSLOT];
#pragma unroll
      for (int i = 0; i < NSLOT; ++i) {
        pr_v[i] = fast_exp2(sc[i] - m_slot[i]);  // masked rows: exp2(-inf) = 0
        l_slot[i] += pr_v[i];
      }
      // publish the (p,p) pairs: per sub-phase VPL consecutive heads of one pass
      if (c * VPL < NV) {
#pragma unroll
        for (int sub = 0; sub < NSUB; ++sub) {
          const uint32_t pa = pw + sub * GK * 8;
          if (VPL == 1) {
            sts64(pa, make_uint2(__float_as_uint(pr_v[sub]), __float_as_uint(pr_v[sub])));
          } else {
#pragma unroll
            for (int i = 0; i < VPL; i += 2)
              sts128(pa + i * 8,
                     make_uint4(__float_as_uint(pr_v[sub * VPL + i]), __float_as_uint(pr_v[sub * VPL + i]),
                                __float_as_uint(pr_v[sub * VPL + (i + 1) % VPL]),
                                __float_as_uint(pr_v[sub * VPL + (i + 1) % VPL])));
          }
        }
      }
    } else {
      // ================================================================ V: accumulate P.V
#pragma unroll
      for (int j = 0; j < R; ++j) {
        u64 v2[NP];
        Chunk<T>::widen(lds128(tile + j * RPP * ROWB), v2);
        const uint32_t pa = pr + j * GT * 8;
        if (GT == 1) {
          uint2 pp = lds64(pa);
          const u64 p2 = pack2(__uint_as_float(pp.x), __uint_as_float(pp.y));
#pragma unroll
          for (int e = 0; e < NP; ++e) o2[0][e] = ffma2(p2, v2[e], o2[0][e]);
        } else {
#pragma unroll
          for (int g = 0; g < GT; g += 2) {
            uint4 pp = lds128(pa + g * 8);
            const u64 pA = pack2(__uint_as_float(pp.x), __uint_as_float(pp.y));
            const u64 pB = pack2(__uint_as_float(pp.z), __uint_as_float(pp.w));
#pragma unroll
            for (int e = 0; e < NP; ++e) {
              o2[g][e] = ffma2(pA, v2[e], o2[g][e]);
              o2[(g + 1) % GT][e] = ffma2(pB, v2[e], o2[(g + 1) % GT][e]);
            }
          }
        }
      }
    }
  }
  cp_async_wait<0>();
  __syncwarp();

  // ---- warp state -> merge area (aliases this warp's ring): o[h][g][D] fp32, then (m,l)[g]
  const uint32_t mo = ring;
  const uint32_t mml = ring + RPP * GT * D * 4;
  {
    // l per head: gather every lane's slot sums through the transpose buffer
#pragma unroll
    for (int i = 0; i < NSLOT; ++i) sts32(xbuf + (lane * NSLOT + i) * 4, __float_as_uint(l_slot[i]));
    __syncwarp();
    float lsum[GT];
#pragma unroll
    for (int g = 0; g < GT; ++g) lsum[g] = 0.f;
#pragma unroll
    for (int L = 0; L < 32; ++L)
#pragma unroll
      for (int i = 0; i < NSLOT; ++i) {
        const int idx = (L % LPR) * VPL + i % VPL;
        if (idx < NV)
          lsum[(i / VPL) * GK + idx % GK] += __uint_as_float(lds32(xbuf + (L * NSLOT + i) * 4));
      }
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      float lo, hi;
#pragma unroll
      for (int e = 0; e < NP; e += 2) {
        float a0, a1, a2, a3;
        unpack2(o2[g][e], a0, a1);
        unpack2(o2[g][e + 1], a2, a3);
        sts128(mo + ((h * GT + g) * D + c * C::VEC + e * 2) * 4,
               make_uint4(__float_as_uint(a0), __float_as_uint(a1), __float_as_uint(a2),
                          __float_as_uint(a3)));
      }
      (void)lo;
      (void)hi;
      if (lane == 0)
        sts64(mml + g * 8, make_uint2(__float_as_uint(m_all[g]), __float_as_uint(lsum[g])));
    }
  }
  __syncthreads();

  // ---- merge the NW warps and write the row (one chunk) or an fp32 partial (several chunks)
  const uint32_t sbase = smem_u32(smem);
  for (int item = threadIdx.x; item < GT * (D / 4); item += NW * 32) {
    const int g = item / (D / 4), dq = item % (D / 4);
    if (g >= g_valid) continue;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w)
      M = fmaxf(M, __uint_as_float(lds32(sbase + w * C::WARP_SMEM + RPP * GT * D * 4 + g * 8)));
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, Lsum = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const uint32_t wb = sbase + w * C::WARP_SMEM;
      uint2 ml = lds64(wb + RPP * GT * D * 4 + g * 8);
      const float sw = fast_exp2(__uint_as_float(ml.x) - M);  // idle warp: exp2(-inf) = 0
      Lsum += sw * __uint_as_float(ml.y);
#pragma unroll
      for (int hh = 0; hh < RPP; ++hh) {
        uint4 v = lds128(wb + ((hh * GT + g) * D + dq * 4) * 4);
        a0 += sw * __uint_as_float(v.x);
        a1 += sw * __uint_as_float(v.y);
        a2 += sw * __uint_as_float(v.z);
        a3 += sw * __uint_as_float(v.w);
      }
    }
    const long long bh = (long long)b * p.H + head0 + g;
    if (n_act == 1) {
      const float inv = 1.f / Lsum;
      store_out4<T, D>(p, b, head0 + g, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
    } else {
      const long long pi = bh * p.max_splits + split;
      *reinterpret_cast<float4*>(p.part_o + pi * D + dq * 4) = make_float4(a0, a1, a2, a3);
      if (dq == 0) p.part_ml[pi] = make_float2(M, Lsum);
    }
  }
  if (n_act == 1 || p.counters == nullptr) return;

  // ---- fused combine: the last of this (sequence, head tile)'s n_act CTAs to arrive merges them.
  // The flag lives in dynamic shared memory (warp 0's probability buffer, idle by now): a static
  // __shared__ variable would shift the dynamic base and break the transpose buffer's alignment.
  const uint32_t flag = sbase + C::RING * TILEB + C::XB_BYTES;  // warp 0's probability buffer
  __syncthreads();  // every thread's partial stores are issued
  if (threadIdx.x == 0) {
    __threadfence();  // release: the CTA's partials (ordered by the barrier) before the count
    int* cnt = p.counters + (long long)b * gridDim.x + blockIdx.x;
    const int prev = atomicAdd(cnt, 1);
    const int last = (prev == n_act - 1);
    if (last) {
      *cnt = 0;         // leave the counter ready for the next call
      __threadfence();  // acquire: the other CTAs' partials after the count
    }
    sts32(flag, (uint32_t)last);
  }
  __syncthreads();
  const bool s_last = lds32(flag) != 0;
  if (!s_last) return;
  for (int item = threadIdx.x; item < GT * (D / 4); item += NW * 32) {
    const int g = item / (D / 4), dq = item % (D / 4);
    if (g >= g_valid) continue;
    const long long pi0 = ((long long)b * p.H + head0 + g) * p.max_splits;
    float M = -INFINITY, Lsum = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    // eight partials' loads in flight at a time, folded in chunk order (the merge is the last thing on a
    // sequence's critical path; see paged_decode_mma.cuh)
    constexpr int MB = 8;
    for (int s0 = 0; s0 < n_act; s0 += MB) {
      float2 ml[MB];
      float4 v[MB];
#pragma unroll
      for (int u = 0; u < MB; ++u) {
        if (s0 + u < n_act) {
          ml[u] = __ldcg(p.part_ml + pi0 + s0 + u);
          v[u] = __ldcg(reinterpret_cast<const float4*>(p.part_o + (pi0 + s0 + u) * D + dq * 4));
        }
      }
#pragma unroll
      for (int u = 0; u < MB; ++u) {
        if (s0 + u >= n_act) break;
        const float Mn = fmaxf(M, ml[u].x);
        const float so = fast_exp2(M - Mn), sn = fast_exp2(ml[u].x - Mn);  // first: exp2(-inf) = 0
        Lsum = Lsum * so + ml[u].y * sn;
        a0 = a0 * so + v[u].x * sn;
        a1 = a1 * so + v[u].y * sn;
        a2 = a2 * so + v[u].z * sn;
        a3 = a3 * so + v[u].w * sn;
        M = Mn;
      }
    }
    const float inv = 1.f / Lsum;
    store_out4<T, D>(p, b, head0 + g, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
  }
}

// Combine kernel: one warp per (sequence, head); merges the active chunks' partials.
// out = sum_s 2^(m_s - M) o_s / sum_s 2^(m_s - M) l_s.  Rows with <= 1 chunk were written by the
// split kernel and are skipped here.
template <typename T, int D>
__global__ void __launch_bounds__(128) paged_decode_combine_kernel(const DecodeParams p, long long rows) {
  constexpr int EPL = D / 32;  // floats per lane (2 or 4)
  const long long bh = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (bh >= rows) return;
  const long long b = bh / p.H;
  long long ctx_ll = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b]
                                  : (long long)reinterpret_cast<const int*>(p.ctx)[b];
  if (ctx_ll > p.max_ctx) ctx_ll = p.max_ctx;
  if (ctx_ll <= p.chunk) return;
  const int n_act = (int)((ctx_ll + p.chunk - 1) / p.chunk);
  const float2* ml = p.part_ml + bh * p.max_splits;
  const float* po = p.part_o + bh * p.max_splits * D + lane * EPL;

  float M = -INFINITY;
  for (int s = lane; s < n_act; s += 32) M = fmaxf(M, ml[s].x);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));

  float acc[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) acc[e] = 0.f;
  float Lsum = 0.f;
  int s = 0;
  for (; s + 4 <= n_act; s += 4) {
    float w[4], v[4][EPL];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float2 x = ml[s + u];
      w[u] = fast_exp2(x.x - M);
      Lsum += w[u] * x.y;
      if (EPL == 4) {
        float4 t = *reinterpret_cast<const float4*>(po + (long long)(s + u) * D);
        v[u][0] = t.x; v[u][1] = t.y; v[u][2 % EPL] = t.z; v[u][3 % EPL] = t.w;
      } else {
        float2 t = *reinterpret_cast<const float2*>(po + (long long)(s + u) * D);
        v[u][0] = t.x; v[u][1] = t.y;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int e = 0; e < EPL; ++e) acc[e] += w[u] * v[u][e];
  }
  for (; s < n_act; ++s) {
    const float2 x = ml[s];
    const float w = fast_exp2(x.x - M);
    Lsum += w * x.y;
#pragma unroll
    for (int e = 0; e < EPL; ++e) acc[e] += w * po[(long long)s * D + e];
  }
  const float inv = 1.f / Lsum;
  const int head = (int)(bh - b * p.H);
  if (EPL == 4) {
    store_out4<T, D>(p, b, head, lane * EPL, acc[0] * inv, acc[1] * inv, acc[2 % EPL] * inv,
                     acc[3 % EPL] * inv);
  } else {
    store_out2<T, D>(p, b, head, lane * EPL, acc[0] * inv, acc[1] * inv);
  }
}

}  // namespace mvl
