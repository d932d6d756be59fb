// allgather.cu — mvl_allgather_heads: the NCCL form of the head gather behind the C ABI.
//
// After head-sharded decode attention rank r holds (batch, heads/N, head_dim); a consumer that needs
// all heads wants (batch, heads, head_dim) on every rank (BASELINE north_star: "an NCCL all-gather
// over NVLink only where the downstream o_proj needs full heads").  ncclAllGather concatenates on the
// outermost dimension, the head dimension is the middle one, so the gather lands in a caller-owned
// staging buffer (N, batch, heads/N, head_dim) and one small kernel lays it out as (batch, heads,
// head_dim).  The fused alternative without a collective is mvl_decode_step_peers_sync.
//
// NCCL is not a link-time dependency of this library: the caller's process already holds the NCCL
// that created `nccl_comm`, and a second copy must not be mixed with it.  ncclAllGather is looked up
// in the libnccl.so.2 that is already loaded (dlopen with RTLD_NOLOAD), falling back to the loader's
// search path when none is.
#include <dlfcn.h>
#include <stdio.h>

#include <mutex>

#include "host.h"
#include "mvl_abi.h"

namespace {

typedef int (*AllGatherFn)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef const char* (*ErrStrFn)(int);

std::mutex g_nccl_mutex;
AllGatherFn g_allgather = nullptr;
ErrStrFn g_errstr = nullptr;
bool g_nccl_tried = false;

AllGatherFn nccl_allgather() {
  std::lock_guard<std::mutex> lock(g_nccl_mutex);
  if (!g_nccl_tried) {
    g_nccl_tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (h) {
      g_allgather = reinterpret_cast<AllGatherFn>(dlsym(h, "ncclAllGather"));
      g_errstr = reinterpret_cast<ErrStrFn>(dlsym(h, "ncclGetErrorString"));
    }
  }
  return g_allgather;
}

// staging (N, B, row16) -> out (B, N, row16), 16-byte chunks; row16 = heads/N * head_dim * es / 16
__global__ void gather_layout_kernel(const uint4* __restrict__ staging, uint4* __restrict__ out, long long batch,
                                     int world, int row16) {
  const long long total = batch * world * row16;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % row16);
    const long long br = i / row16;  // = b * world + r in the output
    const int r = (int)(br % world);
    const long long b = br / world;
    out[i] = staging[((long long)r * batch + b) * row16 + c];
  }
}

}  // namespace

extern "C" int mvl_allgather_heads(void* out_full, void* staging, const void* out_local, int64_t batch,
                                   int heads_local, int head_dim, int elem_dtype, int world_size, void* nccl_comm,
                                   void* stream) {
  const int es = elem_dtype == MVL_DTYPE_F32 ? 4 : (elem_dtype == MVL_DTYPE_F16 || elem_dtype == MVL_DTYPE_BF16) ? 2 : 0;
  if (es == 0) return MVL_ERR_DTYPE;
  if (batch < 0 || heads_local <= 0 || head_dim <= 0 || world_size < 1) return MVL_ERR_SHAPE;
  if (batch == 0) return MVL_OK;
  if (!out_full || !out_local || (world_size > 1 && (!staging || !nccl_comm))) return MVL_ERR_NULL_POINTER;
  const size_t row_bytes = (size_t)heads_local * head_dim * es;
  if (row_bytes % 16 || ((uintptr_t)out_full & 15) || ((uintptr_t)out_local & 15) || ((uintptr_t)staging & 15))
    return MVL_ERR_ALIGNMENT;
  cudaStream_t s = (cudaStream_t)stream;
  if (world_size == 1) {
    if (out_full != out_local) {
      cudaError_t e = cudaMemcpyAsync(out_full, out_local, (size_t)batch * row_bytes, cudaMemcpyDeviceToDevice, s);
      if (e != cudaSuccess) return mvl::fail_cuda((int)e, "cudaMemcpyAsync (single-rank gather)");
    }
    return MVL_OK;
  }
  AllGatherFn ag = nccl_allgather();
  if (!ag) return mvl::fail_text("libnccl.so.2 / ncclAllGather not found in this process", 0);
  const int rc = ag(out_local, staging, (size_t)batch * row_bytes, /*ncclInt8*/ 0, nccl_comm, s);
  if (rc != 0) return mvl::fail_text(g_errstr ? g_errstr(rc) : "ncclAllGather failed", rc);
  const int row16 = (int)(row_bytes / 16);
  const long long total = batch * world_size * row16;
  const unsigned blocks = (unsigned)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
  gather_layout_kernel<<<blocks, 256, 0, s>>>(static_cast<const uint4*>(staging), static_cast<uint4*>(out_full),
                                              batch, world_size, row16);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return mvl::fail_cuda((int)e, "gather_layout launch");
  return MVL_OK;
}
