// prefix_share.cu — device pre-pass that finds sequences whose block tables alias (prefix caching).
//
// The reference's block manager shares physical blocks between sequences with a common prefix
// (reference src/myvllm/engine/block_manager.py:39-44, 67-97), so their block-table rows start with
// the same ids.  Nothing in the reference's kernels uses that; here the decode kernel can score
// several such sequences against one copy of the K/V tiles.  This file
// finds the sharing from the tables alone, on the device, with no host read-back:
//   lcp_kernel    lcpt[b] = tokens sequence b shares with b-1: leading equal table entries x
//                 block_size, capped below both sequences' last token (whose page is never shared)
//   group_kernel  runs of consecutive sequences sharing at least one chunk are cut into sub-groups
//                 of `mm` members; leader[b], gsize[leader], gshare[leader] = tokens shared by all
#include "common.cuh"
#include "host.h"

namespace mvl {

__device__ __forceinline__ long long load_ctx(const void* ctx, int is_i64, long long b, long long cap) {
  long long v = is_i64 ? reinterpret_cast<const long long*>(ctx)[b]
                       : (long long)reinterpret_cast<const int*>(ctx)[b];
  return v > cap ? cap : v;
}

__global__ void __launch_bounds__(512)
lcp_kernel(const int32_t* __restrict__ bt, long long bt_stride, const void* ctx, int ctx_is_i64,
           long long max_ctx, int bs_shift, int* __restrict__ lcpt) {
  const int b = blockIdx.x;
  __shared__ int s_min[16];
  griddep_launch_dependents();  // group_kernel may be set up meanwhile; it waits for this grid before reading
  if (b == 0) {
    if (threadIdx.x == 0) lcpt[0] = 0;
    return;
  }
  const long long ca = load_ctx(ctx, ctx_is_i64, b, max_ctx), cb = load_ctx(ctx, ctx_is_i64, b - 1, max_ctx);
  const long long lim = (ca < cb ? ca : cb) - 1;  // never the last token of either
  if (lim <= 0) {
    if (threadIdx.x == 0) lcpt[b] = 0;
    return;
  }
  const int nblk = (int)((lim + (1 << bs_shift) - 1) >> bs_shift);
  const int32_t* ra = bt + (long long)b * bt_stride;
  const int32_t* rb = ra - bt_stride;
  int first = nblk;  // first mismatching entry seen by this thread
  // eight strided entries of both rows in flight per pass: with one compare per dependent round trip a
  // 128k-token row (8192 entries) took 29 us, a tenth of the decode step it precedes
  for (int j0 = threadIdx.x; j0 < nblk && j0 < first; j0 += 8 * blockDim.x) {
    int a[8], c[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int j = j0 + u * blockDim.x;
      a[u] = j < nblk ? __ldg(ra + j) : 0;
      c[u] = j < nblk ? __ldg(rb + j) : 0;
    }
#pragma unroll
    for (int u = 7; u >= 0; --u)
      if (a[u] != c[u]) first = min(first, j0 + u * (int)blockDim.x);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
  if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = first;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) first = min(first, s_min[w]);
    const long long tok = (long long)first << bs_shift;
    lcpt[b] = (int)(tok < lim ? tok : lim);
  }
}

// One CTA.  meta = leader[B] | gsize[B] | gshare[B] | lcpt[B].
__global__ void __launch_bounds__(256) group_kernel(int* __restrict__ meta, int batch, int chunk, int mm,
                                                    int mm_narrow) {
  int* leader = meta;
  int* gsize = meta + batch;
  int* gshare = meta + 2 * batch;
  const int* lcpt = meta + 3 * batch;
  __shared__ int s_warp[8];
  __shared__ int s_carry;
  griddep_launch_dependents();  // the decode kernels behind wait for this grid before they read `meta`
  griddep_wait();               // lcpt of every sequence is written
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  // run start of b = last j <= b that does not share a whole chunk with its predecessor
  for (int base = 0; base < batch; base += 256) {
    const int b = base + threadIdx.x;
    int v = (b < batch && (b == 0 || lcpt[b] < chunk)) ? b : -1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      if ((threadIdx.x & 31) >= o) v = max(v, u);
    }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = v;
    __syncthreads();
    int pre = s_carry;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) pre = max(pre, s_warp[w]);
    v = max(v, pre);
    if (b < batch) leader[b] = v + ((b - v) / mm) * mm;  // sub-groups of mm inside the run
    __syncthreads();
    if (threadIdx.x == 255) s_carry = v;
    __syncthreads();
  }
  for (int b = threadIdx.x; b < batch; b += 256) {
    int sz = 1, sh = 0x7fffffff;
    if (leader[b] == b) {
      for (int m = 1; m < mm && b + m < batch && leader[b + m] == b; ++m) {
        ++sz;
        sh = min(sh, lcpt[b + m]);
      }
    }
    gsize[b] = sz;
    gshare[b] = sz > 1 ? sh : 0;
  }
  // work list of the wide sub-groups (more than mm_narrow members, at least one whole shared chunk), in
  // batch order; one thread: there are at most batch / (mm_narrow + 1) of them
  __syncthreads();
  if (threadIdx.x == 0) {
    int* wl_b = meta + 4 * batch;
    int* wl_cum = meta + 5 * batch;
    int n = 0, cum = 0;
    for (int b = 0; b < batch; ++b) {
      if (leader[b] == b && gsize[b] > mm_narrow && gshare[b] >= chunk) {
        wl_b[n] = b;
        wl_cum[n] = cum;
        cum += gshare[b] / chunk;
        ++n;
      }
    }
    meta[6 * batch] = n;
    meta[6 * batch + 1] = cum;
  }
}

int launch_prefix_share(const int32_t* bt, long long bt_stride, const void* ctx, int ctx_is_i64,
                        long long batch, long long max_ctx, int bs_shift, int chunk, int mm, int mm_narrow,
                        int* meta, cudaStream_t s) {
  lcp_kernel<<<(unsigned)batch, 512, 0, s>>>(bt, bt_stride, ctx, ctx_is_i64, max_ctx, bs_shift,
                                             meta + 3 * batch);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  // programmatic dependent of lcp_kernel: its launch overlaps that kernel, it waits for it before reading
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(256);
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, group_kernel, meta, (int)batch, chunk, mm, mm_narrow);
}

}  // namespace mvl
