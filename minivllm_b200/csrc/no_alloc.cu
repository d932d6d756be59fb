// no_alloc.cu — mvl_paged_decode_static: paged decode attention for callers that cannot hand us
// scratch memory.
//
// Who needs this.  mvl_paged_decode wants a caller-owned workspace for the split-KV partials.
// Some hosts have no way to produce one:
//   * an engine that gave every free byte of HBM to its KV cache at start-up (the reference sizes
//     its cache that way) has nothing left to allocate from once it is serving;
//   * a foreign-language binding that was only ever handed raw device pointers by its framework
//     owns no allocator, and a cudaMalloc behind the framework's back is not allowed to it;
//   * a latency-critical loop cannot afford cudaMalloc at all, because it synchronises the device.
// For them the partials live in a fixed array inside this library's device image.  The price is
// that every GPU which loads the library carries MVL_STATIC_SCRATCH_BYTES of it until the process
// exits, and that problems whose partials do not fit are refused.
//
// The array below is a __device__ symbol.  In host code its name does not denote a device address;
// the address is looked up with cudaGetSymbolAddress, once per device, while that device is
// current, and the looked-up pointer is what reaches the kernels.  The looked-up address is kept
// for the life of the process: a cudaDeviceReset() unloads the device image and makes it stale,
// so a process that resets a device must not call this entry point on that device afterwards.
#include <stdio.h>

#include <mutex>

#include "host.h"
#include "mvl_abi.h"

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[MVL_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
void* g_scratch_addr[kMaxDevices] = {};  // per device ordinal; nullptr = not looked up yet

thread_local char g_static_err[384] = "";

// Device address of g_static_scratch on the CURRENT device (looked up on first use there).
int scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    snprintf(g_static_err, sizeof(g_static_err), "cudaGetDevice: %s (device %d)",
             cudaGetErrorString(e), dev);
    return MVL_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_scratch_addr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr) {
      snprintf(g_static_err, sizeof(g_static_err), "cudaGetSymbolAddress(static scratch): %s",
               cudaGetErrorString(e));
      return MVL_ERR_CUDA;
    }
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return MVL_OK;
}

}  // namespace

extern "C" {

size_t mvl_static_scratch_bytes(void) { return MVL_STATIC_SCRATCH_BYTES; }

const char* mvl_static_last_error(void) { return g_static_err; }

int mvl_paged_decode_static(void* out, const void* query, const void* k_cache, const void* v_cache,
                            const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                            int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                            int num_heads, int num_kv_heads, int head_dim, int block_size,
                            int64_t num_blocks, float scale, int elem_dtype, int chunk_tokens,
                            void* stream) {
  g_static_err[0] = '\0';
  // Size first: refuse what does not fit before touching the device.
  const size_t need = mvl_paged_decode_workspace_bytes(batch, num_heads, num_kv_heads, head_dim,
                                                       max_num_blocks, block_size, elem_dtype,
                                                       chunk_tokens);
  if (need > (size_t)MVL_STATIC_SCRATCH_BYTES) {
    snprintf(g_static_err, sizeof(g_static_err),
             "split-KV partials need %zu bytes but the built-in scratch holds %zu; pass a larger "
             "chunk_tokens (fewer chunks per sequence), a smaller batch, or use mvl_paged_decode "
             "with your own workspace",
             need, (size_t)MVL_STATIC_SCRATCH_BYTES);
    return MVL_ERR_LIMIT;
  }
  void* scratch = nullptr;
  const int st = scratch_for_current_device(&scratch);
  if (st != MVL_OK) return st;
  // Nothing in the scratch has to start at zero: the combine kernel reads exactly the partials the
  // split kernel of THIS call wrote (the chunks each sequence really has), so leftovers of an
  // earlier call are never read and no clearing pass is launched.
  const int rc = mvl_paged_decode(out, query, k_cache, v_cache, block_tables, context_lens,
                                  ctx_is_i64, batch, max_num_blocks, block_table_stride, num_heads,
                                  num_kv_heads, head_dim, block_size, num_blocks, scale, elem_dtype,
                                  scratch, (size_t)MVL_STATIC_SCRATCH_BYTES, chunk_tokens, 0u,
                                  stream);
  if (rc != MVL_OK) {  // argument check or launch failure of the shared decode path
    const char* cuda = mvl_last_cuda_error();
    snprintf(g_static_err, sizeof(g_static_err), "%s%s%s", mvl_strerror(rc),
             (rc == MVL_ERR_CUDA && cuda && cuda[0]) ? ": " : "",
             (rc == MVL_ERR_CUDA && cuda) ? cuda : "");
  }
  return rc;
}

}  // extern "C"
