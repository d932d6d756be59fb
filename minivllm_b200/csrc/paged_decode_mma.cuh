// paged_decode_mma.cuh — the split-KV decode kernel with the score and P.V products on the tensor
// cores (mma.sync m16n8k16, fp32 accumulation), for 16-bit KV caches.
//
// Why it exists.  BASELINE.json's north star keeps decode attention on the CUDA cores ("a
// bandwidth-bound GEMV").  Measured on B200 that holds only partly for GQA: the FFMA2 kernel spends
// ~800 warp instructions per 8 KB of K/V (G = 4), so (i) under a sustained run the 1000 W power cap
// pulls the SM clock from 1965 to ~1750 MHz and the kernel follows it down from 95 % to ~87 % of
// the HBM peak, (ii) G = 8 (Qwen3-32B shape) is FP32-issue bound at 64 %, and (iii) shared-prefix
// batches that hit L2 are bound the same way.  A GQA group's scores ARE a small GEMM — (G heads x D)
// times (D x tokens) — and with tokens as the M dimension the instruction's N = 8 is exactly the
// largest head tile.  It is the default for 16-bit caches (MVL_DECODE_CUDA_CORES selects the FFMA2 kernel) and is measured against the FFMA2
// kernel in profiles/; DESIGN.md records the deviation and the numbers.
//
// Formulation (per warp, per 16-token tile; T = bf16 or fp16):
//   S^T (16 tokens x 8 heads) = K_tile (16 x D) . Q^T (D x 8)       D/16 x mma, A = K via ldmatrix.x4,
//                                                                  B = q fragments held in registers
//   p = 2^(S^T * scale*log2e - m)   lazily refreshed running max per head, fp32
//   P^T -> (hi, lo) pair of 16-bit values, so the product keeps ~16 mantissa bits of p
//   O^T (D x 8 heads) += V_tile^T (D x 16) . P^T (16 x 8)           D/16 x 2 mma, A = V via
//                                                                  ldmatrix.x4.trans, B = P^T moved
//                                                                  from the accumulator layout with
//                                                                  movmatrix (8x8 transpose)
// Everything else — grid, chunks, per-warp cp.async ring, fused store, fused combine, merge — is
// the FFMA2 kernel's (paged_decode.cuh).  Shared-memory rows are XOR-swizzled by 16-byte chunk
// (chunk ^ (row & 7)) so that ldmatrix reads are bank-conflict free.
#pragma once
#include "paged_decode.cuh"

namespace mvl {

template <typename T> struct MmaType;
template <> struct MmaType<__nv_bfloat16> {
  static __device__ __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  // (lo, hi) fp32 -> packed 16-bit pair, low half = lo; and back to fp32
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
  }
  static __device__ __forceinline__ float lo_f(uint32_t v) { return __uint_as_float(v << 16); }
  static __device__ __forceinline__ float hi_f(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
};
template <> struct MmaType<__half> {
  static __device__ __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
  }
  static __device__ __forceinline__ float lo_f(uint32_t v) {
    return __half2float(__ushort_as_half((unsigned short)(v & 0xffffu)));
  }
  static __device__ __forceinline__ float hi_f(uint32_t v) {
    return __half2float(__ushort_as_half((unsigned short)(v >> 16)));
  }
};

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
// 8x8 transpose of 16-bit elements across the warp: in (row = lane/4, cols 2*(lane%4), +1)
// -> out (row' = old col, cols' = old rows).
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

// NT = column tiles per CTA: 1 (8 columns) normally, 2 (16 columns) when prefix sharing packs up
// to 16/Gp sequences into one CTA.
template <typename T, int D_, int NT_ = 1>
struct MmaCfg {
  static constexpr int NT = NT_;
  static constexpr int NC = 8 * NT_;           // columns = (sequence, head) pairs
  static constexpr int D = D_;
  static constexpr int GT = 8;                 // heads per CTA = N of the instruction
  static constexpr int NW = 4;
  static constexpr int TR = 16;                // tokens per warp tile = M (scores) = K (P.V)
  static constexpr int ROWB = D * (int)sizeof(T);
  static constexpr int CPRW = ROWB / 16;       // 16-byte chunks per row
  static constexpr int RPP = 32 / CPRW;        // rows one cp.async pass of the warp covers
  static constexpr int R = TR / RPP;           // passes per tile
  static constexpr int TILEB = TR * ROWB;
  static constexpr int KS = D / 16;            // k-steps of the score product = m-tiles of P.V
  static constexpr int RING = 3;
  static constexpr int MRG_BYTES = NC * D * 4 + NC * 8;  // merged [column][D] fp32 + (m,l)[column]
  static constexpr int WARP_SMEM = (RING * TILEB > MRG_BYTES ? RING * TILEB : ((MRG_BYTES + 255) / 256) * 256);
  // q of the CTA's columns, staged once through shared memory (rows padded by 16 B so that the 8
  // fragment rows a load touches fall into different banks): one coalesced 16-byte read per chunk
  // instead of every warp gathering its fragments from global memory in 4-byte pieces.
  static constexpr int QROWB = ROWB + 16;
  static constexpr int Q_BYTES = ((NC * QROWB + 255) / 256) * 256;
  static constexpr int SMEM = NW * WARP_SMEM + Q_BYTES;
  static_assert(sizeof(T) == 2, "tensor-core path is for 16-bit caches");
  static_assert(CPRW >= 8 && CPRW <= 16, "head_dim 64 or 128");
};

template <typename T, int D, bool UNIFORM, int NT>
__device__ __forceinline__ void paged_decode_split_mma_body(const DecodeParams& p) {
  using C = MmaCfg<T, D, NT>;
  constexpr int NC = C::NC;
  using M = MmaType<T>;
  constexpr int GT = C::GT, NW = C::NW, TR = C::TR, ROWB = C::ROWB, CPRW = C::CPRW, RPP = C::RPP,
                R = C::R, TILEB = C::TILEB, KS = C::KS, RING = C::RING;
  extern __shared__ __align__(1024) unsigned char smem[];
  // With prefix sharing in its wide form a second kernel (paged_decode_group.cuh) takes the chunks
  // shared by large sub-groups; it depends on nothing this kernel writes, so it may start at once
  // (programmatic dependent launch; a no-op when no dependent launch is configured).
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;  // fragment coordinates: group (row) and thread-in-group
  const int kvh = blockIdx.x / p.n_gt, gt = blockIdx.x % p.n_gt;
  const int b = blockIdx.y, split = blockIdx.z;
  const int head0 = kvh * p.G + gt * GT;
  const int g_valid = min(GT, p.G - gt * GT);

  long long ctx_ll = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b]
                                  : (long long)reinterpret_cast<const int*>(p.ctx)[b];
  if (ctx_ll > p.max_ctx) ctx_ll = p.max_ctx;
  const int ctx = (int)ctx_ll;
  if (ctx <= 0) {
    if (split == 0) {
      if (!p.beside_group) griddep_wait();
      for (int i = threadIdx.x; i < g_valid * (D / 4); i += NW * 32)
        store_out4<T, D>(p, b, head0 + i / (D / 4), (i % (D / 4)) * 4, 0.f, 0.f, 0.f, 0.f);
    }
    return;
  }
  const int start = split * p.chunk;
  if (start >= ctx) return;
  const int end = min(ctx, start + p.chunk);
  const int n_act = (ctx + p.chunk - 1) / p.chunk;

  // Programmatic dependent launch: this grid may have been started while the kernel in front of it
  // on the stream is still finishing.  What that kernel cannot have touched — context lengths, block
  // tables and the K/V rows of tokens older than the step's new one — is fetched before
  // griddep_wait(); everything it may have produced (q, the new token's K/V row, the prefix-sharing
  // groups) and every global write comes after.
  const bool holds_last = ctx - 1 >= start && ctx - 1 < end;
  const bool prefetch_early = p.overlap_prev && p.share_meta == nullptr && !holds_last;
  if (!prefetch_early && !p.beside_group) griddep_wait();

  if (p.k_new != nullptr && holds_last) {  // fused store (see paged_decode.cuh)
    const long long slot = p.slot_is_i64 ? reinterpret_cast<const long long*>(p.slot_map)[b]
                                         : (long long)reinterpret_cast<const int*>(p.slot_map)[b];
    if (slot >= 0 && slot < p.num_slots) {
      if (threadIdx.x < 2 * CPRW) {
        const int isv = threadIdx.x / CPRW, ch = threadIdx.x % CPRW;
        const char* src = reinterpret_cast<const char*>(isv ? p.v_new : p.k_new) +
                          b * (isv ? p.v_new_stride : p.k_new_stride) + (long long)kvh * ROWB + ch * 16;
        char* dst = reinterpret_cast<char*>(const_cast<void*>(isv ? p.vc : p.kc)) +
                    (slot * p.KVH + kvh) * ROWB + ch * 16;
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
      }
      __threadfence();
    }
    __syncthreads();
  }

  // ---- prefix sharing: if every member of b's sub-group shares this whole chunk, the leader's
  // CTA scores all of them (columns = member x head) and the other members' CTAs have nothing to do
  int nmem = 1;
  if (p.share_meta != nullptr) {
    const int nb = gridDim.y;
    const int lead = p.share_meta[b];
    const int sz = p.share_meta[nb + lead];
    if (sz > 1 && start + p.chunk <= p.share_meta[2 * nb + lead]) {
      if (b != lead || sz > p.mm_narrow) return;  // larger sub-groups: the group kernel's chunk
      nmem = sz;
    }
  }
  const int gp_mask = (1 << p.gp_shift) - 1;

  const uint32_t ring = smem_u32(smem) + warp * C::WARP_SMEM;

  // ---- this warp's tiles and the cp.async ring (same scheme as the FFMA2 kernel)
  const int ntiles = (end - start + TR - 1) / TR;
  const int my_nt = (warp < ntiles) ? (ntiles - warp + NW - 1) / NW : 0;
  const uint32_t row_pitch = (uint32_t)p.KVH * ROWB;
  const int bs_mask = (1 << p.bs_shift) - 1;
  const int32_t* bt_row = p.bt + (long long)b * p.bt_stride;
  const int ch = lane % CPRW, hrow = lane / CPRW;  // this lane's chunk and row inside a pass
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
        const int rr = j * RPP + hrow;  // row inside the tile
        const uint32_t dst = dst0 + rr * ROWB + ((ch ^ (rr & 7)) << 4);  // chunk swizzle
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
  if (prefetch_early) griddep_wait();

  // ---- q as B fragments of the score product: n = column nt*8+g = (member, head), k = d.
  // Staged through shared memory: thread i copies 16-byte chunk i of the CTA's q rows (coalesced; also
  // the only q traffic when q lives in host memory), then every lane picks its fragments from there.
  const uint32_t qs = smem_u32(smem) + NW * C::WARP_SMEM;
  for (int i = threadIdx.x; i < NC * CPRW; i += NW * 32) {
    const int col = i / CPRW, c16 = i % CPRW;
    const int cm = col >> p.gp_shift, chd = col & gp_mask;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (cm < nmem && chd < g_valid)
      v = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(p.q) +
                                          ((long long)(b + cm) * p.H + head0 + chd) * ROWB + c16 * 16);
    sts128(qs + col * C::QROWB + c16 * 16, v);
  }
  __syncthreads();
  uint32_t qb[NT][KS][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const uint32_t qrow = qs + (nt * 8 + g) * C::QROWB + 4 * t;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      qb[nt][ks][0] = lds32(qrow + ks * 32);
      qb[nt][ks][1] = lds32(qrow + ks * 32 + 16);
    }
  }

  // ---- state per column tile: O^T accumulators (d x 8 columns), running max / sum of this thread's
  // two columns (2t, 2t+1), P^T B fragments
  float o[NT][KS][4];
  float m_a[NT], m_b[NT], l_a[NT], l_b[NT];
  uint32_t pb_hi[NT][2], pb_lo[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
    for (int mt = 0; mt < KS; ++mt) o[nt][mt][0] = o[nt][mt][1] = o[nt][mt][2] = o[nt][mt][3] = 0.f;
    m_a[nt] = m_b[nt] = -INFINITY;
    l_a[nt] = l_b[nt] = 0.f;
    pb_hi[nt][0] = pb_hi[nt][1] = pb_lo[nt][0] = pb_lo[nt][1] = 0u;
  }

  // ldmatrix lane addressing: matrix mi = lane/8, row lane%8
  const int mi = lane >> 3, mr = lane & 7;
  const int k_tok = (mi & 1) * 8 + mr, k_chk = mi >> 1;   // K: matrices (tok lo|hi) x (chunk 2ks|2ks+1)
  const int v_tok = (mi >> 1) * 8 + mr, v_chk = mi & 1;   // V^T: (chunk 2mt|2mt+1) x (tok lo|hi)
  const uint32_t k_off = k_tok * ROWB, v_off = v_tok * ROWB;

  int slot_c = 0;
#pragma unroll 1
  for (int n = 0; n < 2 * my_nt; ++n) {
    cp_async_wait<RING - 2>();
    __syncwarp();
    issue(n + RING - 1, slot_c == 0 ? RING - 1 : slot_c - 1);
    const uint32_t tile = ring + slot_c * TILEB;
    slot_c = (slot_c + 1 == RING) ? 0 : slot_c + 1;
    if (!(n & 1)) {
      // ================================================================ scores S^T = K . Q^T
      const int tstart = start + (warp + (n >> 1) * NW) * TR;
      float s[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t a[4];
        ldmatrix_x4(a, tile + k_off + (((2 * ks + k_chk) ^ (k_tok & 7)) << 4));
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) M::mma(s[nt], a, qb[nt][ks][0], qb[nt][ks][1]);
      }
      // s[.][0]: (token g, column 2t)  [1]: (token g, column 2t+1)  [2], [3]: token g+8
      const bool cut = tstart + TR > end;
      bool need = false;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) s[nt][i] *= p.qscale;
        if (cut) {
          if (tstart + g >= end) s[nt][0] = s[nt][1] = -INFINITY;
          if (tstart + g + 8 >= end) s[nt][2] = s[nt][3] = -INFINITY;
        }
        need |= fmaxf(s[nt][0], s[nt][2]) > m_a[nt] + 8.f || fmaxf(s[nt][1], s[nt][3]) > m_b[nt] + 8.f;
      }
      if (__any_sync(0xffffffffu, need)) {  // rare: refresh the running maxima, rescale the state
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          float xa = fmaxf(s[nt][0], s[nt][2]), xb = fmaxf(s[nt][1], s[nt][3]);
#pragma unroll
          for (int off = 4; off < 32; off <<= 1) {
            xa = fmaxf(xa, __shfl_xor_sync(0xffffffffu, xa, off));
            xb = fmaxf(xb, __shfl_xor_sync(0xffffffffu, xb, off));
          }
          const float na = fmaxf(m_a[nt], xa), nb = fmaxf(m_b[nt], xb);
          const float fa = fast_exp2(m_a[nt] - na), fb = fast_exp2(m_b[nt] - nb);  // -inf -> 0
          l_a[nt] *= fa;
          l_b[nt] *= fb;
#pragma unroll
          for (int mt = 0; mt < KS; ++mt) {
            o[nt][mt][0] *= fa;
            o[nt][mt][2] *= fa;
            o[nt][mt][1] *= fb;
            o[nt][mt][3] *= fb;
          }
          m_a[nt] = na;
          m_b[nt] = nb;
        }
      }
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        float pr[4];
        pr[0] = fast_exp2(s[nt][0] - m_a[nt]);
        pr[1] = fast_exp2(s[nt][1] - m_b[nt]);
        pr[2] = fast_exp2(s[nt][2] - m_a[nt]);
        pr[3] = fast_exp2(s[nt][3] - m_b[nt]);
        // split p into a 16-bit value and a 16-bit correction; sums use what the products will use
        const uint32_t h0 = M::pack(pr[0], pr[1]), h1 = M::pack(pr[2], pr[3]);
        const float r0 = pr[0] - M::lo_f(h0), r1 = pr[1] - M::hi_f(h0);
        const float r2 = pr[2] - M::lo_f(h1), r3 = pr[3] - M::hi_f(h1);
        const uint32_t c0 = M::pack(r0, r1), c1 = M::pack(r2, r3);
        l_a[nt] += (M::lo_f(h0) + M::lo_f(c0)) + (M::lo_f(h1) + M::lo_f(c1));
        l_b[nt] += (M::hi_f(h0) + M::hi_f(c0)) + (M::hi_f(h1) + M::hi_f(c1));
        // (token, columns) 8x8 blocks -> (column, tokens): the B fragments of the P.V product
        pb_hi[nt][0] = movmatrix_trans(h0);
        pb_hi[nt][1] = movmatrix_trans(h1);
        pb_lo[nt][0] = movmatrix_trans(c0);
        pb_lo[nt][1] = movmatrix_trans(c1);
      }
    } else {
      // ================================================================ O^T += V^T . P^T
#pragma unroll
      for (int mt = 0; mt < KS; ++mt) {
        uint32_t a[4];
        ldmatrix_x4_trans(a, tile + v_off + (((2 * mt + v_chk) ^ (v_tok & 7)) << 4));
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          M::mma(o[nt][mt], a, pb_hi[nt][0], pb_hi[nt][1]);
          M::mma(o[nt][mt], a, pb_lo[nt][0], pb_lo[nt][1]);
        }
      }
    }
  }
  cp_async_wait<0>();
  __syncwarp();

  // ---- warp state -> merge area (aliases the ring): o[column][D] fp32, then (m, l)[column]
  const uint32_t mo = ring, mml = ring + NC * D * 4;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
      l_a[nt] += __shfl_xor_sync(0xffffffffu, l_a[nt], off);
      l_b[nt] += __shfl_xor_sync(0xffffffffu, l_b[nt], off);
    }
    const int ca = nt * 8 + 2 * t, cb = ca + 1;
#pragma unroll
    for (int mt = 0; mt < KS; ++mt) {
      // o[nt][mt][0]: (d = 16mt+g, column 2t)  [1]: column 2t+1  [2],[3]: d + 8
      sts32(mo + (ca * D + 16 * mt + g) * 4, __float_as_uint(o[nt][mt][0]));
      sts32(mo + (cb * D + 16 * mt + g) * 4, __float_as_uint(o[nt][mt][1]));
      sts32(mo + (ca * D + 16 * mt + g + 8) * 4, __float_as_uint(o[nt][mt][2]));
      sts32(mo + (cb * D + 16 * mt + g + 8) * 4, __float_as_uint(o[nt][mt][3]));
    }
    if (g == 0) {
      sts64(mml + ca * 8, make_uint2(__float_as_uint(m_a[nt]), __float_as_uint(l_a[nt])));
      sts64(mml + cb * 8, make_uint2(__float_as_uint(m_b[nt]), __float_as_uint(l_b[nt])));
    }
  }
  __syncthreads();

  // ---- merge the warps; per column (member, head): write the row (one chunk) or a partial
  const uint32_t sbase = smem_u32(smem);
  auto member_chunks = [&](int cm) -> int {  // number of chunks sequence b+cm really has
    if (cm == 0) return n_act;
    long long c = p.ctx_is_i64 ? reinterpret_cast<const long long*>(p.ctx)[b + cm]
                               : (long long)reinterpret_cast<const int*>(p.ctx)[b + cm];
    if (c > p.max_ctx) c = p.max_ctx;
    return (int)((c + p.chunk - 1) / p.chunk);
  };
  for (int item = threadIdx.x; item < NC * (D / 4); item += NW * 32) {
    const int col = item / (D / 4), dq = item % (D / 4);
    const int cm = col >> p.gp_shift, hd = col & gp_mask;
    if (cm >= nmem || hd >= g_valid) continue;
    float Mx = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w)
      Mx = fmaxf(Mx, __uint_as_float(lds32(sbase + w * C::WARP_SMEM + NC * D * 4 + col * 8)));
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, Lsum = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const uint32_t wb = sbase + w * C::WARP_SMEM;
      const uint2 ml = lds64(wb + NC * D * 4 + col * 8);
      const float sw = fast_exp2(__uint_as_float(ml.x) - Mx);
      Lsum += sw * __uint_as_float(ml.y);
      const uint4 v = lds128(wb + (col * D + dq * 4) * 4);
      a0 += sw * __uint_as_float(v.x);
      a1 += sw * __uint_as_float(v.y);
      a2 += sw * __uint_as_float(v.z);
      a3 += sw * __uint_as_float(v.w);
    }
    if (member_chunks(cm) == 1) {
      const float inv = 1.f / Lsum;
      store_out4<T, D>(p, b + cm, head0 + hd, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
    } else {
      const long long pi = ((long long)(b + cm) * p.H + head0 + hd) * p.max_splits + split;
      *reinterpret_cast<float4*>(p.part_o + pi * D + dq * 4) = make_float4(a0, a1, a2, a3);
      if (dq == 0) p.part_ml[pi] = make_float2(Mx, Lsum);
    }
  }
  if (p.counters == nullptr) return;
  if (nmem == 1 && n_act == 1) return;

  // ---- fused combine, per member: the last of a sequence's chunk CTAs to arrive merges its partials
  const uint32_t flag = sbase + C::WARP_SMEM - 16;  // last bytes of warp 0's region (unused by the merge)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    uint32_t last_mask = 0;
    for (int cm = 0; cm < nmem; ++cm) {
      const int na = member_chunks(cm);
      if (na == 1) continue;
      int* cnt = p.counters + (long long)(b + cm) * gridDim.x + blockIdx.x;
      if (atomicAdd(cnt, 1) == na - 1) {
        *cnt = 0;
        last_mask |= 1u << cm;
      }
    }
    if (last_mask) __threadfence();
    sts32(flag, last_mask);
  }
  __syncthreads();
  const uint32_t last_mask = lds32(flag);
  if (last_mask == 0) return;
  for (int item = threadIdx.x; item < NC * (D / 4); item += NW * 32) {
    const int col = item / (D / 4), dq = item % (D / 4);
    const int cm = col >> p.gp_shift, hd = col & gp_mask;
    if (cm >= nmem || hd >= g_valid || !((last_mask >> cm) & 1u)) continue;
    const int na = member_chunks(cm);
    const long long pi0 = ((long long)(b + cm) * p.H + head0 + hd) * p.max_splits;
    float Mx = -INFINITY, Lsum = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    // eight partials' loads in flight at a time (the merge is the last thing on a sequence's critical
    // path: with one dependent L2 round trip per partial, and still with four per round trip, many-chunk
    // splits paid for it — C5's 32 chunks: 8 round trips of ~1.5 us behind the last byte streamed),
    // folded in chunk order so that the result does not depend on the batching
    constexpr int MB = 8;
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
    store_out4<T, D>(p, b + cm, head0 + hd, dq * 4, a0 * inv, a1 * inv, a2 * inv, a3 * inv);
  }
}

template <typename T, int D, bool UNIFORM, int NT>
__global__ void __launch_bounds__(128, NT == 1 ? 4 : 2) paged_decode_split_mma_kernel(const DecodeParams p) {
  paged_decode_split_mma_body<T, D, UNIFORM, NT>(p);  // every exit from it is CTA-uniform
  // Beside the group kernel: this grid's completion must imply that grid's, and ONE CTA waiting for it is
  // enough for that (the last in launch order).  If every CTA waited, the thousands that have nothing to do
  // would sit in all the slots the group CTAs free and keep the working ones out until that grid is done.
  if (p.beside_group && blockIdx.x == gridDim.x - 1 && blockIdx.y == gridDim.y - 1 && blockIdx.z == gridDim.z - 1)
    griddep_wait();
  // ---- head sharding with peer stores (peer_sync.cu): the last CTA of the whole grid to get here —
  // CTAs whose chunk lies past their sequence's end included — announces the step to every rank
  if (p.sync_local != nullptr) {
    // A CTA that had nothing to do never reached the body's griddepcontrol.wait; with a programmatic
    // dependency it may be running while the PREVIOUS launch of this kernel is still counting arrivals on
    // the same counter (seen at N = 8: a step announced early, the next one never, every rank in its
    // time-out).  No arrival before the kernel in front has completed.
    griddep_wait();
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();  // this CTA's rows (peer stores included) before its arrival
      if (atomicAdd(p.sync_local + 9, 1u) == gridDim.x * gridDim.y * gridDim.z - 1u) {
        p.sync_local[9] = 0u;
        const unsigned step = p.sync_local[8] + 1u;
        p.sync_local[8] = step;
        __threadfence_system();
        for (int r = 0; r < p.n_out; ++r)
          asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p.sync_peers[r] + p.my_rank), "r"(step) : "memory");
      }
    }
  }
}

}  // namespace mvl
