// paged_decode_stream.cuh — the tensor-core decode kernel as a persistent, evenly balanced stream.
//
// Why it exists.  The split kernel (paged_decode_mma.cuh) cuts every sequence into chunks of a
// host-chosen size and gives each (sequence, kv head, chunk) its own CTA.  That is at its best when
// there are many more chunks than resident CTAs.  When there are not — one kv head per rank under
// 8-way head sharding, a batch of a few long sequences, one layer of a small model — the launch is a
// wave and a bit: C3 on an 8-way shard is ~700 chunks of 1 MB for 592 resident CTAs, so the last
// hundred CTAs run alone on a mostly idle chip (117 us measured against 93 us of streaming time).
//
// Here the whole launch is ONE stream of 64-token rows, ordered (sequence, kv head, position), cut
// into equal pieces, one piece per resident CTA (grid = 4 CTAs x SMs, every CTA gets the same
// number of rows +-1 whatever the batch looks like).  A piece may start in the middle of one
// sequence and end in the middle of another: a CTA walks the segments of its piece one after the
// other, and for each segment does exactly what the split kernel does for a chunk.  Where a
// (sequence, kv head) unit lies entirely inside one piece the row is written directly; where it is
// spread over several pieces each writes a partial (at most two per CTA: the unit that was already
// under way when its piece began and the one still under way when it ends) and the last to arrive
// merges them in piece order — deterministic whatever the arrival order.
//
// What equal pieces do not buy (measured, scripts/diag_stream_times.py): CTAs with equal work do not finish
// together.  On C3's 8-way shard the median CTA is done at 87 us and the last at ~100, the same blocks early
// and late in every launch; sizing the pieces by a probe's per-SM speed or by each block's own speed in the
// previous launch leaves that spread where it is.  Handing the last quarter of the
// stream out dynamically (items claimed with an atomic
// counter, halving in size) was built and measured: the fine items it needs near the end turn every
// long sequence into dozens of partials whose merge sits on the critical path (0.11 -> 0.20 ms on that
// shard), so the cut stays static.
//
// The piece boundaries come from the context lengths, which live on the device: every CTA scans them
// itself at start (batch <= 1024 values from L2, ~1 us), so there is no pre-pass launch, no host
// read-back, and the grid depends only on host-known sizes (graph-capturable like the split kernel).
#pragma once
#include "paged_decode_mma.cuh"

namespace mvl {

constexpr int kStreamMaxBatch = 1024;   // context lengths scanned per CTA; larger batches use the split kernel
constexpr int kStreamRow = 64;          // tokens per stream row = 4 warps x 16-token tiles
constexpr int kStreamMinRows = 2;       // smallest piece: below this the partials outweigh the K/V read

template <typename T, int D>
struct StreamCfg {
  using M = MmaCfg<T, D, 1>;
  static constexpr int CUM_BYTES = ((kStreamMaxBatch + 1) * 4 + 255) / 256 * 256;
  static constexpr int MISC_BYTES = 256;  // warp totals of the scan, combine flag
  static constexpr int SMEM = M::SMEM + CUM_BYTES + MISC_BYTES;
  static constexpr int PART_FLOATS = 8 * D;  // one partial = 8 columns x D fp32 (+ 8 (m, l) pairs)
};

// Bytes of partials for `ctas` pieces (two partial slots per piece).
inline size_t stream_partial_bytes(int ctas, int D) {
  return (size_t)ctas * 2 * (8 * (size_t)D * sizeof(float) + 8 * sizeof(float2));
}

template <typename T, int D, bool UNIFORM>
__global__ void __launch_bounds__(128, 4) paged_decode_stream_mma_kernel(const DecodeParams p, const int batch) {
  using C = MmaCfg<T, D, 1>;
  using S = StreamCfg<T, D>;
  using M = MmaType<T>;
  constexpr int GT = C::GT, NW = C::NW, TR = C::TR, ROWB = C::ROWB, CPRW = C::CPRW, RPP = C::RPP,
                R = C::R, TILEB = C::TILEB, KS = C::KS, RING = C::RING;
  extern __shared__ __align__(1024) unsigned char smem[];
  // (C::SMEM = the four warps' rings + the q staging area of paged_decode_mma.cuh)
  int* cum = reinterpret_cast<int*>(smem + C::SMEM);                 // cum[b] = rows of sequences < b
  int* misc = reinterpret_cast<int*>(smem + C::SMEM + S::CUM_BYTES);  // [0..3] warp totals, [8] flag

  // Programmatic dependent launch (see paged_decode_mma.cuh): the scan of the context lengths and the
  // search for this CTA's piece run while the kernel in front is still finishing.
  griddep_launch_dependents();
  auto stamp = [&](int which) {
    if (p.dbg_times != nullptr && threadIdx.x == 0) {
      unsigned long long tnow;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tnow));
      p.dbg_times[(size_t)blockIdx.x * 5 + which] = tnow;
      if (which == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.dbg_times[(size_t)blockIdx.x * 5 + 4] = smid;
      }
    }
  };
  stamp(0);
  bool first_tile_stamped = false;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int X = p.KVH * p.n_gt;  // units per sequence

  auto ctx_of = [&](int b) -> int {
    long long c = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b]
                               : (long long)reinterpret_cast<const int*>(p.ctx)[b];
    if (c > p.max_ctx) c = p.max_ctx;
    return c > 0 ? (int)c : 0;
  };

  // ---- exclusive scan of rows per sequence (every CTA does its own; batch <= kStreamMaxBatch)
  {
    const int per = (batch + NW * 32 - 1) / (NW * 32);  // <= 8
    const int b0 = threadIdx.x * per;
    int sum = 0;
    for (int j = 0; j < per; ++j) {
      const int b = b0 + j;
      if (b < batch) sum += (ctx_of(b) + kStreamRow - 1) / kStreamRow;
    }
    int incl = sum;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += v;
    }
    if (lane == 31) misc[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < warp; ++w) base += misc[w];
    int run = base + incl - sum;
    for (int j = 0; j < per; ++j) {
      const int b = b0 + j;
      if (b < batch) {
        cum[b] = run;
        run += (ctx_of(b) + kStreamRow - 1) / kStreamRow;
      }
    }
    if (threadIdx.x == NW * 32 - 1) cum[batch] = run;
    __syncthreads();
  }

  const long long total = (long long)cum[batch] * X;
  long long Tll = (total + gridDim.x - 1) / gridDim.x;
  if (Tll < kStreamMinRows) Tll = kStreamMinRows;
  const int Tp = (int)Tll;                          // rows per piece
  const long long piece0 = (long long)blockIdx.x * Tp;
  // a CTA past the end of the stream has no piece (it still takes part in the completion signal)
  const long long piece1 = piece0 >= total ? piece0 : ((piece0 + Tp < total) ? piece0 + Tp : total);

  // first sequence of the piece: largest b with X * cum[b] <= piece0
  int b = 0;
  if (piece0 < total) {
    int lo = 0, hi = batch;  // invariant: X*cum[lo] <= piece0 < X*cum[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if ((long long)cum[mid] * X <= piece0) lo = mid; else hi = mid;
    }
    b = lo;
  }

  const uint32_t sbase = smem_u32(smem);
  const uint32_t ring = sbase + warp * C::WARP_SMEM;
  const uint32_t row_pitch = (uint32_t)p.KVH * ROWB;
  const int bs_mask = (1 << p.bs_shift) - 1;
  const int ch = lane % CPRW, hrow = lane / CPRW;
  const int mi = lane >> 3, mr = lane & 7;
  const int k_tok = (mi & 1) * 8 + mr, k_chk = mi >> 1;
  const int v_tok = (mi >> 1) * 8 + mr, v_chk = mi & 1;
  const uint32_t k_off = k_tok * ROWB, v_off = v_tok * ROWB;

  griddep_wait();
  stamp(1);
  // rows with no context: zeros (the split kernel's convention), spread over the CTAs
  for (int bz = blockIdx.x; bz < batch; bz += gridDim.x) {
    if (cum[bz + 1] == cum[bz])
      for (int i = threadIdx.x; i < p.H * (D / 4); i += NW * 32)
        store_out4<T, D>(p, bz, i / (D / 4), (i % (D / 4)) * 4, 0.f, 0.f, 0.f, 0.f);
  }

  long long pos = piece0;
#pragma unroll 1
  while (pos < piece1) {
    const int tb = cum[b + 1] - cum[b];  // rows of sequence b (per unit)
    if (tb == 0) { ++b; continue; }
    const long long seq0 = (long long)cum[b] * X;
    const int off = (int)(pos - seq0);
    if (off >= tb * X) { ++b; continue; }
    const int x = off / tb, r0 = off - x * tb;
    const long long left = piece1 - pos;
    const int r1 = (left < (long long)(tb - r0)) ? r0 + (int)left : tb;
    const bool first_seg = (pos == piece0);
    // where this unit's rows lie in the stream, hence which pieces hold parts of it
    const long long us = seq0 + (long long)x * tb, ue = us + tb;
    const int it_first = (int)(us / Tp), it_last = (int)((ue - 1) / Tp);
    const int nparts = it_last - it_first + 1;
    pos += r1 - r0;

    const int kvh = x / p.n_gt, gt = x % p.n_gt;
    const int head0 = kvh * p.G + gt * GT;
    const int g_valid = min(GT, p.G - gt * GT);
    const int ctx = ctx_of(b);
    const int start = r0 * kStreamRow;
    const int end = min(ctx, r1 * kStreamRow);

    if (p.k_new != nullptr && ctx - 1 >= start && ctx - 1 < end) {  // fused store (paged_decode.cuh)
      const long long slot = p.slot_is_i64 ? reinterpret_cast<const long long*>(p.slot_map)[b]
                                           : (long long)reinterpret_cast<const int*>(p.slot_map)[b];
      if (slot >= 0 && slot < p.num_slots) {
        if (threadIdx.x < 2 * CPRW) {
          const int isv = threadIdx.x / CPRW, c16 = threadIdx.x % CPRW;
          const char* src = reinterpret_cast<const char*>(isv ? p.v_new : p.k_new) +
                            b * (isv ? p.v_new_stride : p.k_new_stride) + (long long)kvh * ROWB + c16 * 16;
          char* dst = reinterpret_cast<char*>(const_cast<void*>(isv ? p.vc : p.kc)) +
                      (slot * p.KVH + kvh) * ROWB + c16 * 16;
          *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
        }
        __threadfence();
      }
      __syncthreads();
    }

    // ---- q as B fragments (n = head g, k = d), staged through shared memory as in the split kernel.
    // The previous segment's fragment reads of the staging area are behind that segment's barriers.
    const uint32_t qs = sbase + NW * C::WARP_SMEM;
    for (int i = threadIdx.x; i < 8 * CPRW; i += NW * 32) {
      const int col = i / CPRW, c16 = i % CPRW;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (col < g_valid)
        v = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(p.q) +
                                            ((long long)b * p.H + head0 + col) * ROWB + c16 * 16);
      sts128(qs + col * C::QROWB + c16 * 16, v);
    }
    __syncthreads();
    uint32_t qb[KS][2];
    {
      const uint32_t qrow = qs + g * C::QROWB + 4 * t;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        qb[ks][0] = lds32(qrow + ks * 32);
        qb[ks][1] = lds32(qrow + ks * 32 + 16);
      }
    }

    const int ntiles = (end - start + TR - 1) / TR;
    const int my_nt = (warp < ntiles) ? (ntiles - warp + NW - 1) / NW : 0;
    const int32_t* bt_row = p.bt + (long long)b * p.bt_stride;
    const char* kbase = reinterpret_cast<const char*>(p.kc) + (long long)kvh * ROWB + ch * 16;
    const char* vbase = reinterpret_cast<const char*>(p.vc) + (long long)kvh * ROWB + ch * 16;

    auto table_entry = [&](int i) -> uint32_t {
      return (UNIFORM && i < my_nt)
                 ? (uint32_t)__ldg(bt_row + ((start + (warp + i * NW) * TR) >> p.bs_shift))
                 : 0u;
    };
    uint32_t blk_a = table_entry(0), blk_b = table_entry(1);

    auto issue = [&](int n, int slot) {
      if (n < 2 * my_nt) {
        const int tstart = start + (warp + (n >> 1) * NW) * TR;
        const char* base = (n & 1) ? vbase : kbase;
        const uint32_t dst0 = ring + slot * TILEB;
#pragma unroll
        for (int j = 0; j < R; ++j) {
          const int rr = j * RPP + hrow;
          const uint32_t dst = dst0 + rr * ROWB + ((ch ^ (rr & 7)) << 4);
          const int tok = tstart + rr;
          const bool valid = tok < end;
          if (UNIFORM) {
            const uint32_t row0 = (blk_a << p.bs_shift) + (tstart & bs_mask);
            cp_async16(dst, row_ptr(base, row0 + rr, row_pitch), valid ? 16u : 0u);
          } else {
            const int tt = valid ? tok : end - 1;
            const uint32_t blk = (uint32_t)__ldg(bt_row + (tt >> p.bs_shift));
            cp_async16(dst, row_ptr(base, (blk << p.bs_shift) + (tt & bs_mask), row_pitch), valid ? 16u : 0u);
          }
        }
      }
      cp_async_commit();
      if (UNIFORM && (n & 1)) {
        blk_a = blk_b;
        blk_b = table_entry((n >> 1) + 2);
      }
    };
#pragma unroll 1
    for (int n = 0; n < RING - 1; ++n) issue(n, n);

    float o[KS][4];
#pragma unroll
    for (int mt = 0; mt < KS; ++mt) o[mt][0] = o[mt][1] = o[mt][2] = o[mt][3] = 0.f;
    float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.f, l_b = 0.f;
    uint32_t pb_hi[2] = {0u, 0u}, pb_lo[2] = {0u, 0u};

    int slot_c = 0;
#pragma unroll 1
    for (int n = 0; n < 2 * my_nt; ++n) {
      cp_async_wait<RING - 2>();
      __syncwarp();
      if (!first_tile_stamped) {
        stamp(2);
        first_tile_stamped = true;
      }
      issue(n + RING - 1, slot_c == 0 ? RING - 1 : slot_c - 1);
      const uint32_t tile = ring + slot_c * TILEB;
      slot_c = (slot_c + 1 == RING) ? 0 : slot_c + 1;
      if (!(n & 1)) {
        const int tstart = start + (warp + (n >> 1) * NW) * TR;
        float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t a[4];
          ldmatrix_x4(a, tile + k_off + (((2 * ks + k_chk) ^ (k_tok & 7)) << 4));
          M::mma(s, a, qb[ks][0], qb[ks][1]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) s[i] *= p.qscale;
        if (tstart + TR > end) {
          if (tstart + g >= end) s[0] = s[1] = -INFINITY;
          if (tstart + g + 8 >= end) s[2] = s[3] = -INFINITY;
        }
        const bool need = fmaxf(s[0], s[2]) > m_a + 8.f || fmaxf(s[1], s[3]) > m_b + 8.f;
        if (__any_sync(0xffffffffu, need)) {
          float xa = fmaxf(s[0], s[2]), xb = fmaxf(s[1], s[3]);
#pragma unroll
          for (int sh = 4; sh < 32; sh <<= 1) {
            xa = fmaxf(xa, __shfl_xor_sync(0xffffffffu, xa, sh));
            xb = fmaxf(xb, __shfl_xor_sync(0xffffffffu, xb, sh));
          }
          const float na = fmaxf(m_a, xa), nb = fmaxf(m_b, xb);
          const float fa = fast_exp2(m_a - na), fb = fast_exp2(m_b - nb);
          l_a *= fa;
          l_b *= fb;
#pragma unroll
          for (int mt = 0; mt < KS; ++mt) {
            o[mt][0] *= fa;
            o[mt][2] *= fa;
            o[mt][1] *= fb;
            o[mt][3] *= fb;
          }
          m_a = na;
          m_b = nb;
        }
        float pr[4];
        pr[0] = fast_exp2(s[0] - m_a);
        pr[1] = fast_exp2(s[1] - m_b);
        pr[2] = fast_exp2(s[2] - m_a);
        pr[3] = fast_exp2(s[3] - m_b);
        const uint32_t h0 = M::pack(pr[0], pr[1]), h1 = M::pack(pr[2], pr[3]);
        const float q0 = pr[0] - M::lo_f(h0), q1 = pr[1] - M::hi_f(h0);
        const float q2 = pr[2] - M::lo_f(h1), q3 = pr[3] - M::hi_f(h1);
        const uint32_t c0 = M::pack(q0, q1), c1 = M::pack(q2, q3);
        l_a += (M::lo_f(h0) + M::lo_f(c0)) + (M::lo_f(h1) + M::lo_f(c1));
        l_b += (M::hi_f(h0) + M::hi_f(c0)) + (M::hi_f(h1) + M::hi_f(c1));
        pb_hi[0] = movmatrix_trans(h0);
        pb_hi[1] = movmatrix_trans(h1);
        pb_lo[0] = movmatrix_trans(c0);
        pb_lo[1] = movmatrix_trans(c1);
      } else {
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          uint32_t a[4];
          ldmatrix_x4_trans(a, tile + v_off + (((2 * mt + v_chk) ^ (v_tok & 7)) << 4));
          M::mma(o[mt], a, pb_hi[0], pb_hi[1]);
          M::mma(o[mt], a, pb_lo[0], pb_lo[1]);
        }
      }
    }
    cp_async_wait<0>();
    __syncwarp();

    // ---- warp state -> merge area (aliases the ring)
    const uint32_t mo = ring, mml = ring + 8 * D * 4;
#pragma unroll
    for (int sh = 4; sh < 32; sh <<= 1) {
      l_a += __shfl_xor_sync(0xffffffffu, l_a, sh);
      l_b += __shfl_xor_sync(0xffffffffu, l_b, sh);
    }
    {
      const int ca = 2 * t, cb = ca + 1;
#pragma unroll
      for (int mt = 0; mt < KS; ++mt) {
        sts32(mo + (ca * D + 16 * mt + g) * 4, __float_as_uint(o[mt][0]));
        sts32(mo + (cb * D + 16 * mt + g) * 4, __float_as_uint(o[mt][1]));
        sts32(mo + (ca * D + 16 * mt + g + 8) * 4, __float_as_uint(o[mt][2]));
        sts32(mo + (cb * D + 16 * mt + g + 8) * 4, __float_as_uint(o[mt][3]));
      }
      if (g == 0) {
        sts64(mml + ca * 8, make_uint2(__float_as_uint(m_a), __float_as_uint(l_a)));
        sts64(mml + cb * 8, make_uint2(__float_as_uint(m_b), __float_as_uint(l_b)));
      }
    }
    __syncthreads();

    // ---- merge the warps; per head: the row (unit inside one piece) or this piece's partial
    const long long my_slot = (long long)blockIdx.x * 2 + (first_seg ? 0 : 1);
    for (int item = threadIdx.x; item < 8 * (D / 4); item += NW * 32) {
      const int col = item / (D / 4), dq = item % (D / 4);
      if (col >= g_valid) continue;
      float Mx = -INFINITY;
#pragma unroll
      for (int w = 0; w < NW; ++w)
        Mx = fmaxf(Mx, __uint_as_float(lds32(sbase + w * C::WARP_SMEM + 8 * D * 4 + col * 8)));
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, Lsum = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        const uint32_t wb = sbase + w * C::WARP_SMEM;
        const uint2 ml = lds64(wb + 8 * D * 4 + col * 8);
        const float sw = fast_exp2(__uint_as_float(ml.x) - Mx);
        Lsum += sw * __uint_as_float(ml.y);
        const uint4 v = lds128(wb + (col * D + dq * 4) * 4);
        a0 += sw * __uint_as_float(v.x);
        a1 += sw * __uint_as_float(v.y);
        a2 += sw * __uint_as_float(v.z);
        a3 += sw * __uint_as_float(v.w);
      }
      if (nparts == 1) {
        const float inv = 1.f / Lsum;
        store_out4<T, D>(p, b, head0 + col, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
      } else {
        const long long pi = my_slot * 8 + col;
        *reinterpret_cast<float4*>(p.part_o + pi * D + dq * 4) = make_float4(a0, a1, a2, a3);
        if (dq == 0) p.part_ml[pi] = make_float2(Mx, Lsum);
      }
    }
    __syncthreads();  // the merge area is the ring: the next segment's loads must wait for these reads
    if (nparts == 1) continue;

    // ---- the last of the unit's pieces to arrive merges them, in piece order
    if (threadIdx.x == 0) {
      __threadfence();
      int* cnt = p.counters + (long long)b * X + x;
      int last = 0;
      if (atomicAdd(cnt, 1) == nparts - 1) {
        *cnt = 0;
        last = 1;
        __threadfence();
      }
      misc[8] = last;
    }
    __syncthreads();
    const int is_last = misc[8];
    __syncthreads();  // misc[8] may be rewritten by the next segment
    if (!is_last) continue;
    for (int item = threadIdx.x; item < 8 * (D / 4); item += NW * 32) {
      const int col = item / (D / 4), dq = item % (D / 4);
      if (col >= g_valid) continue;
      float Mx = -INFINITY, Lsum = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      for (int it = it_first; it <= it_last; ++it) {
        // piece `it` holds this unit as its first segment unless the unit begins inside it
        const long long sl = (long long)it * 2 + ((us <= (long long)it * Tp) ? 0 : 1);
        const long long pi = sl * 8 + col;
        const float2 ml = __ldcg(p.part_ml + pi);
        const float4 v = __ldcg(reinterpret_cast<const float4*>(p.part_o + pi * D + dq * 4));
        const float Mn = fmaxf(Mx, ml.x);
        const float so = fast_exp2(Mx - Mn), sn = fast_exp2(ml.x - Mn);
        Lsum = Lsum * so + ml.y * sn;
        a0 = a0 * so + v.x * sn;
        a1 = a1 * so + v.y * sn;
        a2 = a2 * so + v.z * sn;
        a3 = a3 * so + v.w * sn;
        Mx = Mn;
      }
      const float inv = 1.f / Lsum;
      store_out4<T, D>(p, b, head0 + col, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
    }
  }

  stamp(3);
  // ---- head sharding with peer stores: the launch's last CTA announces the step to every rank
  if (p.sync_local != nullptr) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();  // this CTA's rows (peer stores included) before its arrival
      if (atomicAdd(p.sync_local + 9, 1u) == gridDim.x - 1) {
        p.sync_local[9] = 0u;
        const unsigned step = p.sync_local[8] + 1u;
        p.sync_local[8] = step;
        __threadfence_system();  // one fence, then the announcements as relaxed stores (no round trip each)
        for (int r = 0; r < p.n_out; ++r)
          asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p.sync_peers[r] + p.my_rank), "r"(step) : "memory");
      }
    }
  }
}

}  // namespace mvl
