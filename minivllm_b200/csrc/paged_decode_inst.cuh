// paged_decode_inst.cuh — launcher bodies, included once per element type with MVL_T defined.
#include <mutex>

#include "host.h"
#include "paged_decode_mma.cuh"
#include "paged_decode_group.cuh"
#include "paged_decode_stream.cuh"

namespace mvl {

// Launch with programmatic stream serialisation allowed: the kernel may start while its predecessor
// on the stream finishes, and orders itself with griddepcontrol.wait.  Only for kernels that do.
template <typename Kern, typename... Args>
static int launch_pdl(bool pdl, Kern kern, dim3 grid, int block, size_t smem, cudaStream_t s, Args... args) {
  if (!pdl) {
    kern<<<grid, block, smem, s>>>(args...);
    return (int)cudaGetLastError();
  }
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, kern, args...);
}

template <typename T, int D, int GT, bool UNIFORM, bool TMA>
static int launch_split_u(const DecodeParams& p, const CUtensorMap* maps, dim3 grid, cudaStream_t s) {
  using C = DecodeCfg<T, D, GT>;
  auto kern = paged_decode_split_kernel<T, D, GT, UNIFORM, TMA>;
  // Opt in to > 48 KB dynamic shared memory once per device.
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  } else if (dev >= 64) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return (int)e;
  }
  static const CUtensorMap kNoMap = {};
  kern<<<grid, C::NW * 32, C::SMEM, s>>>(p, TMA ? maps[0] : kNoMap, TMA ? maps[1] : kNoMap);
  return (int)cudaGetLastError();
}

template <typename T, int D, int GT>
static int launch_split_one(const DecodeParams& p, const CUtensorMap* maps, dim3 grid, cudaStream_t s) {
  const bool uniform = ((1 << p.bs_shift) % DecodeCfg<T, D, GT>::TR) == 0;
  if (uniform && maps) return launch_split_u<T, D, GT, true, true>(p, maps, grid, s);
  return uniform ? launch_split_u<T, D, GT, true, false>(p, nullptr, grid, s)
                 : launch_split_u<T, D, GT, false, false>(p, nullptr, grid, s);
}

template <typename T, int D>
static int launch_split_d(const DecodeParams& p, const CUtensorMap* maps, int gt, dim3 grid,
                          cudaStream_t s) {
  switch (gt) {
    case 1: return launch_split_one<T, D, 1>(p, maps, grid, s);
    case 2: return launch_split_one<T, D, 2>(p, maps, grid, s);
    case 4: return launch_split_one<T, D, 4>(p, maps, grid, s);
    case 8: return launch_split_one<T, D, 8>(p, maps, grid, s);
  }
  return -1000;
}

template <>
int launch_decode_split<MVL_T>(const DecodeParams& p, const CUtensorMap* maps, int head_dim, int gt,
                               dim3 grid, cudaStream_t s) {
  switch (head_dim) {
    case 64: return launch_split_d<MVL_T, 64>(p, maps, gt, grid, s);
    case 128: return launch_split_d<MVL_T, 128>(p, maps, gt, grid, s);
  }
  return -1000;
}

#ifdef MVL_T_IS_16BIT
template <typename T, int D, bool UNIFORM, int NT>
static int launch_mma_nt(const DecodeParams& p, dim3 grid, cudaStream_t s) {
  using C = MmaCfg<T, D, NT>;
  auto kern = paged_decode_split_mma_kernel<T, D, UNIFORM, NT>;
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) configured[dev] = true;
  }
  return launch_pdl(p.overlap_prev != 0 || p.beside_group != 0, kern, grid, C::NW * 32, C::SMEM, s, p);
}

// One column tile (8 columns).  The kernel is written for NT column tiles; NT = 2 (16 columns, up
// to 16/Gp sequences per CTA under prefix sharing) was measured on C5 and lost to NT = 1: 205
// registers halve the occupancy (0.442 ms vs 0.420 ms), and batches without sharing pay 8 %.
template <typename T, int D, bool UNIFORM>
static int launch_mma_u(const DecodeParams& p, dim3 grid, cudaStream_t s) {
  return launch_mma_nt<T, D, UNIFORM, 1>(p, grid, s);
}

template <>
int launch_decode_split_mma<MVL_T>(const DecodeParams& p, int head_dim, dim3 grid, cudaStream_t s) {
  const bool uniform = ((1 << p.bs_shift) % 16) == 0;
  switch (head_dim) {
    case 64:
      return uniform ? launch_mma_u<MVL_T, 64, true>(p, grid, s) : launch_mma_u<MVL_T, 64, false>(p, grid, s);
    case 128:
      return uniform ? launch_mma_u<MVL_T, 128, true>(p, grid, s) : launch_mma_u<MVL_T, 128, false>(p, grid, s);
  }
  return -1000;
}
#endif

#ifdef MVL_T_IS_16BIT
template <typename T, int D, bool UNIFORM>
static int launch_stream_u(const DecodeParams& p, int batch, int ctas, cudaStream_t s) {
  using S = StreamCfg<T, D>;
  auto kern = paged_decode_stream_mma_kernel<T, D, UNIFORM>;
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::SMEM);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) configured[dev] = true;
  }
  return launch_pdl(p.overlap_prev != 0, kern, dim3((unsigned)ctas), 128, S::SMEM, s, p, batch);
}

template <>
int launch_decode_stream_mma<MVL_T>(const DecodeParams& p, int head_dim, int batch, int ctas, cudaStream_t s) {
  const bool uniform = ((1 << p.bs_shift) % 16) == 0;
  switch (head_dim) {
    case 64:
      return uniform ? launch_stream_u<MVL_T, 64, true>(p, batch, ctas, s)
                     : launch_stream_u<MVL_T, 64, false>(p, batch, ctas, s);
    case 128:
      return uniform ? launch_stream_u<MVL_T, 128, true>(p, batch, ctas, s)
                     : launch_stream_u<MVL_T, 128, false>(p, batch, ctas, s);
  }
  return -1000;
}
#endif

#ifdef MVL_T_IS_16BIT
template <>
int launch_decode_group_mma<MVL_T>(const DecodeParams& p, int head_dim, dim3 grid, cudaStream_t s) {
  static bool configured[64][2] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64) return (int)cudaErrorInvalidDevice;
  const int di = head_dim == 64 ? 0 : 1;
  if (head_dim != 64 && head_dim != 128) return -1000;
  if (!configured[dev][di]) {
    e = di == 0 ? cudaFuncSetAttribute(paged_decode_group_mma_kernel<MVL_T, 64>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GroupCfg<MVL_T, 64>::SMEM)
                : cudaFuncSetAttribute(paged_decode_group_mma_kernel<MVL_T, 128>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GroupCfg<MVL_T, 128>::SMEM);
    if (e != cudaSuccess) return (int)e;
    configured[dev][di] = true;
  }
  // programmatic dependent launch: may start while the split kernel (launched just before on this
  // stream) is still running
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(128);
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (di == 0) {
    cfg.dynamicSmemBytes = GroupCfg<MVL_T, 64>::SMEM;
    e = cudaLaunchKernelEx(&cfg, paged_decode_group_mma_kernel<MVL_T, 64>, p);
  } else {
    cfg.dynamicSmemBytes = GroupCfg<MVL_T, 128>::SMEM;
    e = cudaLaunchKernelEx(&cfg, paged_decode_group_mma_kernel<MVL_T, 128>, p);
  }
  return (int)e;
}
#endif

template <>
int launch_decode_combine<MVL_T>(const DecodeParams& p, int head_dim, long long rows,
                                 cudaStream_t s) {
  const unsigned blocks = (unsigned)((rows + 3) / 4);
  switch (head_dim) {
    case 64: paged_decode_combine_kernel<MVL_T, 64><<<blocks, 128, 0, s>>>(p, rows); break;
    case 128: paged_decode_combine_kernel<MVL_T, 128><<<blocks, 128, 0, s>>>(p, rows); break;
    default: return -1000;
  }
  return (int)cudaGetLastError();
}

template <>
int decode_tile_rows<MVL_T>(int head_dim, int gt) {
  switch (head_dim) {
    case 64: return gt >= 8 ? DecodeCfg<MVL_T, 64, 8>::TR : DecodeCfg<MVL_T, 64, 4>::TR;
    case 128: return gt >= 8 ? DecodeCfg<MVL_T, 128, 8>::TR : DecodeCfg<MVL_T, 128, 4>::TR;
  }
  return -1;
}

}  // namespace mvl
