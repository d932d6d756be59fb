// paged_decode_group.cuh — prefix sharing, wide form: one CTA scores up to 4 x (8 / Gp) sequences of a
// sub-group against ONE copy of the K/V tiles of a chunk they all share.
//
// The tensor-core decode kernel (paged_decode_mma.cuh) packs 8 / Gp sequences into the 8 columns of
// its MMAs and gives each of its 4 warps different TOKEN tiles; a group of 8 sequences with 4 heads
// per KV head therefore still reads every shared tile four times (once per pair of sequences), and
// ncu shows C5 at the L2 -> SM bandwidth cap (4.5 GB moved for 1.9 GB of unique bytes, 11.5 TB/s).
// Here the 4 warps of a CTA hold different MEMBERS (warp w: sequences w * mm .. w * mm + mm - 1 of the
// sub-group, mm = 8 / Gp) and every warp walks ALL token tiles of the chunk: warp w still fetches
// tiles w, w + 4, ... into its own cp.async ring, but after a CTA barrier every warp reads all four
// rings, so a tile crosses L2 -> SM once per 4 * mm sequences.  Rounds of 4 tiles (64 tokens), K and V
// of a round in flight together (4 ring slots per warp), two barriers per round.
// Launched right after the split kernel on the same grid as its programmatic dependent (it runs
// beside it: the two touch disjoint chunks), only when the device pre-pass (prefix_share.cu) can form
// sub-groups larger than mm; the split kernel leaves those chunks alone (DecodeParams::wide).
// Each warp owns its columns outright, so there is no cross-warp merge: the CTA writes its members'
// partials; the fused combine is the split kernel's (per-member arrival counters).  One warp adds the
// tiles four warps split otherwise, and P is rounded to 16 bits once (no correction term): results
// agree with the unshared run to rounding, not bit for bit (MVL_DECODE_SHARE_NARROW keeps that).
// Measured on C5 (B200): 0.398 ms against 0.416 ms for the narrow form — the L2 -> SM traffic of the
// shared chunks drops 4x, but with four sequences' worth of mma.sync per byte the chunk becomes
// issue / tensor-pipe bound (ncu: tensor pipe 44 %, 12 warps per SM); the next step for this case is
// the tcgen05 formulation (K as the M = 128 operand, the group's 32 query rows as N).
#pragma once
#include "paged_decode_mma.cuh"

namespace mvl {

__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename T, int D_>
struct GroupCfg {
  static constexpr int D = D_;
  static constexpr int NW = 4, TR = 16, RING = 4;
  static constexpr int ROWB = D * (int)sizeof(T);
  static constexpr int CPRW = ROWB / 16;
  static constexpr int RPP = 32 / CPRW;
  static constexpr int R = TR / RPP;
  static constexpr int TILEB = TR * ROWB;
  static constexpr int KS = D / 16;
  static constexpr int MRG_BYTES = 8 * D * 4 + 8 * 8;  // this warp's [column][D] fp32 + (m, l)[column]
  static constexpr int WARP_SMEM = RING * TILEB > MRG_BYTES ? RING * TILEB : ((MRG_BYTES + 255) / 256) * 256;
  static constexpr int SMEM = NW * WARP_SMEM;
  static_assert(sizeof(T) == 2, "tensor-core path is for 16-bit caches");
};

template <typename T, int D>
__global__ void __launch_bounds__(128, 3) paged_decode_group_mma_kernel(const DecodeParams p) {
  using C = GroupCfg<T, D>;
  using M = MmaType<T>;
  constexpr int NW = C::NW, TR = C::TR, ROWB = C::ROWB, CPRW = C::CPRW, RPP = C::RPP, R = C::R,
                TILEB = C::TILEB, KS = C::KS;
  extern __shared__ __align__(1024) unsigned char smem[];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int kvh = blockIdx.x / p.n_gt, gt = blockIdx.x % p.n_gt;
  const int b = blockIdx.y, split = blockIdx.z;
  const int head0 = kvh * p.G + gt * 8;
  const int g_valid = min(8, p.G - gt * 8);

  // ---- is this (sequence, chunk) the leader's CTA of a wide, fully shared chunk?
  // Launched as a programmatic dependent of the split kernel and running beside it: nothing here reads
  // what that kernel writes (shared chunks never hold a sequence's newest token; partials and arrival
  // counters are per (sequence, chunk) and atomic).  Every exit first waits for the split kernel's
  // grid (griddepcontrol.wait), so that this kernel's completion implies that one's.
  const int nb = gridDim.y;
  if (p.share_meta[b] != b) { grid_dep_wait(); return; }
  const int nmem = p.share_meta[nb + b];
  const int start = split * p.chunk;
  if (nmem <= p.mm_narrow || start + p.chunk > p.share_meta[2 * nb + b]) { grid_dep_wait(); return; }
  // a fully shared chunk is a whole chunk of every member: tokens [start, start + chunk), none masked
  const int gp_mask = (1 << p.gp_shift) - 1;

  const uint32_t sbase = smem_u32(smem);
  const uint32_t ring = sbase + warp * C::WARP_SMEM;

  // ---- this warp's columns: member w * mm + (column >> gp_shift), head column & gp_mask
  const int col = g, cm = warp * p.mm_narrow + (col >> p.gp_shift), chd = col & gp_mask;
  const bool col_ok = cm < nmem && chd < g_valid;
  const bool warp_live = warp * p.mm_narrow < nmem;  // warp-uniform
  uint32_t qb[KS][2];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    qb[ks][0] = qb[ks][1] = 0u;
    if (col_ok) {
      const T* qrow = reinterpret_cast<const T*>(p.q) + ((long long)(b + cm) * p.H + head0 + chd) * D + ks * 16 + 2 * t;
      qb[ks][0] = *reinterpret_cast<const uint32_t*>(qrow);
      qb[ks][1] = *reinterpret_cast<const uint32_t*>(qrow + 8);
    }
  }

  // ---- loads: warp w fetches tiles w, w + 4, ... (K and V of a round into slots 2 * (round & 1), + 1)
  const int rounds = p.chunk / (NW * TR);  // the chunk is a multiple of 64 tokens (abi.cu, make_plan)
  const uint32_t row_pitch = (uint32_t)p.KVH * ROWB;
  const int bs_mask = (1 << p.bs_shift) - 1;
  const int32_t* bt_row = p.bt + (long long)b * p.bt_stride;  // the members' tables agree on these pages
  const int ch = lane % CPRW, hrow = lane / CPRW;
  const char* kbase = reinterpret_cast<const char*>(p.kc) + (long long)kvh * ROWB + ch * 16;
  const char* vbase = reinterpret_cast<const char*>(p.vc) + (long long)kvh * ROWB + ch * 16;
  auto issue_round = [&](int i) {
    if (i < rounds) {
      const int tstart = start + (warp + i * NW) * TR;
      const uint32_t blk = (uint32_t)__ldg(bt_row + (tstart >> p.bs_shift));
      const uint32_t row0 = (blk << p.bs_shift) + (tstart & bs_mask);  // a 16-token tile lies in one page
      const uint32_t dst0 = ring + (i & 1) * 2 * TILEB;
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const int rr = j * RPP + hrow;
        const uint32_t off = rr * ROWB + ((ch ^ (rr & 7)) << 4);  // chunk swizzle, as in the split kernel
        cp_async16(dst0 + off, row_ptr(kbase, row0 + rr, row_pitch), 16u);
        cp_async16(dst0 + TILEB + off, row_ptr(vbase, row0 + rr, row_pitch), 16u);
      }
    }
    cp_async_commit();
  };
  issue_round(0);
  issue_round(1);

  // ---- state of this warp's 8 columns: O^T accumulators (d x 8), running max / sum of this thread's
  // two columns (2t, 2t+1)
  float o[KS][4];
#pragma unroll
  for (int mt = 0; mt < KS; ++mt) o[mt][0] = o[mt][1] = o[mt][2] = o[mt][3] = 0.f;
  float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.f, l_b = 0.f;

  const int mi = lane >> 3, mr = lane & 7;
  const int k_tok = (mi & 1) * 8 + mr, k_chk = mi >> 1;
  const int v_tok = (mi >> 1) * 8 + mr, v_chk = mi & 1;
  const uint32_t k_off = k_tok * ROWB, v_off = v_tok * ROWB;

#pragma unroll 1
  for (int i = 0; i < rounds; ++i) {
    cp_async_wait<1>();  // this warp's tiles of round i have landed (round i + 1 may be in flight)
    __syncthreads();     // ... and so have the other warps'
    if (warp_live) {
#pragma unroll
      for (int ww = 0; ww < NW; ++ww) {
        const uint32_t ktile = sbase + ww * C::WARP_SMEM + (i & 1) * 2 * TILEB, vtile = ktile + TILEB;
        // ============================================================== scores S^T = K . Q^T
        float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t a[4];
          ldmatrix_x4(a, ktile + k_off + (((2 * ks + k_chk) ^ (k_tok & 7)) << 4));
          M::mma(s, a, qb[ks][0], qb[ks][1]);
        }
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) s[q4] *= p.qscale;
        const bool need = fmaxf(s[0], s[2]) > m_a + 8.f || fmaxf(s[1], s[3]) > m_b + 8.f;
        if (__any_sync(0xffffffffu, need)) {  // rare: refresh the running maxima, rescale the state
          float xa = fmaxf(s[0], s[2]), xb = fmaxf(s[1], s[3]);
#pragma unroll
          for (int off = 4; off < 32; off <<= 1) {
            xa = fmaxf(xa, __shfl_xor_sync(0xffffffffu, xa, off));
            xb = fmaxf(xb, __shfl_xor_sync(0xffffffffu, xb, off));
          }
          const float na = fmaxf(m_a, xa), nbx = fmaxf(m_b, xb);
          const float fa = fast_exp2(m_a - na), fb = fast_exp2(m_b - nbx);  // -inf -> 0
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
          m_b = nbx;
        }
        float pr[4];
        pr[0] = fast_exp2(s[0] - m_a);
        pr[1] = fast_exp2(s[1] - m_b);
        pr[2] = fast_exp2(s[2] - m_a);
        pr[3] = fast_exp2(s[3] - m_b);
        // P as one 16-bit value (the split kernel carries a 16-bit correction as well; here the four
        // warps' worth of MMAs per byte makes the tensor pipe the limit, and one rounding of p is what
        // the prefill kernels and FlashAttention use); the sums use what the products use
        const uint32_t h0 = M::pack(pr[0], pr[1]), h1 = M::pack(pr[2], pr[3]);
        l_a += M::lo_f(h0) + M::lo_f(h1);
        l_b += M::hi_f(h0) + M::hi_f(h1);
        const uint32_t ph0 = movmatrix_trans(h0), ph1 = movmatrix_trans(h1);
        // ============================================================== O^T += V^T . P^T
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          uint32_t a[4];
          ldmatrix_x4_trans(a, vtile + v_off + (((2 * mt + v_chk) ^ (v_tok & 7)) << 4));
          M::mma(o[mt], a, ph0, ph1);
        }
      }
    }
    __syncthreads();       // every warp is done with round i's slots
    issue_round(i + 2);    // ... which round i + 2 reuses
  }
  cp_async_wait<0>();
  __syncthreads();

  // ---- this warp's state -> its merge area (aliases its ring): o[column][D] fp32, then (m, l)[column]
  const uint32_t mo = ring, mml = ring + 8 * D * 4;
#pragma unroll
  for (int off = 4; off < 32; off <<= 1) {
    l_a += __shfl_xor_sync(0xffffffffu, l_a, off);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, off);
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

  // ---- every (member, head) column of the CTA: its partial of this chunk (a member of a shared chunk
  // always has further chunks of its own, so there is never a finished row to write here)
  auto member_chunks = [&](int m) -> int {
    long long c = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b + m]
                               : (long long)reinterpret_cast<const int*>(p.ctx)[b + m];
    if (c > p.max_ctx) c = p.max_ctx;
    return (int)((c + p.chunk - 1) / p.chunk);
  };
  for (int item = threadIdx.x; item < NW * 8 * (D / 4); item += NW * 32) {
    const int gcol = item / (D / 4), dq = item % (D / 4);
    const int w = gcol >> 3, c = gcol & 7;
    const int m = w * p.mm_narrow + (c >> p.gp_shift), hd = c & gp_mask;
    if (m >= nmem || hd >= g_valid) continue;
    const uint32_t wb = sbase + w * C::WARP_SMEM;
    const uint2 ml = lds64(wb + 8 * D * 4 + c * 8);
    const uint4 v = lds128(wb + (c * D + dq * 4) * 4);
    const long long pi = ((long long)(b + m) * p.H + head0 + hd) * p.max_splits + split;
    *reinterpret_cast<float4*>(p.part_o + pi * D + dq * 4) =
        make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w));
    if (dq == 0) p.part_ml[pi] = make_float2(__uint_as_float(ml.x), __uint_as_float(ml.y));
  }
  if (p.counters == nullptr) { grid_dep_wait(); return; }

  // ---- fused combine, per member: the last of a sequence's chunk CTAs to arrive merges its partials
  const uint32_t flag = sbase + C::WARP_SMEM - 16;  // last bytes of warp 0's region (past its merge area)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    uint32_t last_mask = 0;
    for (int m = 0; m < nmem; ++m) {
      const int na = member_chunks(m);
      int* cnt = p.counters + (long long)(b + m) * gridDim.x + blockIdx.x;
      if (atomicAdd(cnt, 1) == na - 1) {
        *cnt = 0;
        last_mask |= 1u << m;
      }
    }
    if (last_mask) __threadfence();
    sts32(flag, last_mask);
  }
  __syncthreads();
  const uint32_t last_mask = lds32(flag);
  if (last_mask == 0) { grid_dep_wait(); return; }
  for (int item = threadIdx.x; item < NW * 8 * (D / 4); item += NW * 32) {
    const int gcol = item / (D / 4), dq = item % (D / 4);
    const int w = gcol >> 3, c = gcol & 7;
    const int m = w * p.mm_narrow + (c >> p.gp_shift), hd = c & gp_mask;
    if (m >= nmem || hd >= g_valid || !((last_mask >> m) & 1u)) continue;
    const int na = member_chunks(m);
    const long long pi0 = ((long long)(b + m) * p.H + head0 + hd) * p.max_splits;
    float Mx = -INFINITY, Lsum = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    constexpr int MB = 8;  // partials' loads in flight at a time; folded in chunk order
    for (int s0 = 0; s0 < na; s0 += MB) {
      float2 ml[MB];
      float4 v[MB];
#pragma unroll
      for (int u = 0; u < MB; ++u) {
        if (s0 + u < na) {
          ml[u] = __ldcg(p.part_ml + pi0 + s0 + u);
          v[u] = __ldcg(reinterpret_cast<const float4*>(p.part_o + (pi0 + s0 + u) * D + dq * 4));
        }
      }
#pragma unroll
      for (int u = 0; u < MB; ++u) {
        if (s0 + u >= na) break;
        const float Mn = fmaxf(Mx, ml[u].x);
        const float so = fast_exp2(Mx - Mn), sn = fast_exp2(ml[u].x - Mn);
        Lsum = Lsum * so + ml[u].y * sn;
        a0 = a0 * so + v[u].x * sn;
        a1 = a1 * so + v[u].y * sn;
        a2 = a2 * so + v[u].z * sn;
        a3 = a3 * so + v[u].w * sn;
        Mx = Mn;
      }
    }
    const float inv = 1.f / Lsum;
    store_out4<T, D>(p, b + m, head0 + hd, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
  }
  grid_dep_wait();
}

}  // namespace mvl
