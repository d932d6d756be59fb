// host.h — internal (non-ABI) declarations shared by the translation units of libmvl_b200.
#pragma once
#include <cuda_runtime.h>

#include "paged_decode.cuh"

namespace mvl {

int launch_store_kvcache(const void* key, const void* value, void* k_cache, void* v_cache,
                         const void* slot_mapping, int slot_is_i64, long long num_tokens,
                         int chunks_per_row, long long key_stride16, long long value_stride16,
                         long long num_slots, int num_sms, cudaStream_t stream);

// Prefix-sharing pre-pass (prefix_share.cu): fills meta[4*batch] = leader | gsize | gshare | lcpt.
// meta[6 * batch + 2] = leader | gsize | gshare | lcpt | wl_b | wl_cum | (wl_n, wl_total): the last three
// list the sub-groups of more than `mm_narrow` members with their shared whole chunks (work list of the
// tcgen05 group kernel).
int launch_prefix_share(const int32_t* bt, long long bt_stride, const void* ctx, int ctx_is_i64,
                        long long batch, long long max_ctx, int bs_shift, int chunk, int mm, int mm_narrow,
                        int* meta, cudaStream_t s);
// Prefix sharing, wide form on tcgen05 (paged_decode_group_tc5.cu): persistent, one CTA per SM, launched
// behind the split kernel as its programmatic dependent.  Returns 0, -1000, or an mvl_status (text recorded).
int launch_decode_group_tc5(const DecodeParams& p, int head_dim, long long batch, int block_size,
                            long long num_blocks, int num_sms, int dtype, cudaStream_t s);

// abi.cu: record an error text for mvl_last_cuda_error() and return MVL_ERR_CUDA.
int fail_text(const char* where, int code);
int fail_cuda(int cuda_error, const char* where);

// peer_sync.cu: announce this rank's step to all ranks / wait until all ranks announced mine.
int launch_peer_signal(unsigned* local, unsigned* const* peers, int n, int my_rank, cudaStream_t s);
int launch_peer_wait(unsigned* local, int n, unsigned long long timeout_ns, cudaStream_t s);

// Launchers, one translation unit per element type (keeps nvcc parallel and fast).
// Return cudaSuccess (0) or the cudaError_t of the failed call; -1000 if (D, GT) is not built.
// maps: {K cache, V cache} tensor maps for the TMA load path, or nullptr for the cp.async path.
template <typename T>
int launch_decode_split(const DecodeParams& p, const CUtensorMap* maps, int head_dim, int gt, dim3 grid,
                        cudaStream_t s);
// Tensor-core variant (16-bit element types only): 8 heads per CTA, 16-token warp tiles.
template <typename T>
int launch_decode_split_mma(const DecodeParams& p, int head_dim, dim3 grid, cudaStream_t s);
// Persistent, evenly balanced form of the tensor-core kernel (paged_decode_stream.cuh): `ctas`
// resident CTAs share one stream of 64-token rows; partials at p.part_o / p.part_ml, two per CTA.
template <typename T>
int launch_decode_stream_mma(const DecodeParams& p, int head_dim, int batch, int ctas, cudaStream_t s);
// Largest batch / partial bytes of that kernel (host-side planning).
constexpr int kStreamBatchLimit = 1024;
inline size_t stream_workspace_partials(int ctas, int head_dim) {
  return (size_t)ctas * 2 * (8 * (size_t)head_dim * sizeof(float) + 8 * sizeof(float2));
}
// Prefix sharing, wide form (paged_decode_group.cuh): same grid as the split kernel, after it.
template <typename T>
int launch_decode_group_mma(const DecodeParams& p, int head_dim, dim3 grid, cudaStream_t s);
template <typename T>
int launch_decode_combine(const DecodeParams& p, int head_dim, long long rows, cudaStream_t s);
// Rows of one warp tile for (T, head_dim, gt): the chunk must be a multiple of 4 warps x this.
template <typename T>
int decode_tile_rows(int head_dim, int gt);

}  // namespace mvl
