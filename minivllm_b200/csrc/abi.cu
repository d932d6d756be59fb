// abi.cu — the extern "C" surface of libmvl_b200.so (declared in include/mvl_abi.h):
// argument validation -> status codes, split-KV planning, kernel dispatch.  No allocation, no
// synchronisation, no host reads of device memory.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "host.h"
#include "prefill.h"
#include "mvl_abi.h"

namespace {

thread_local char g_cuda_err[256] = "";

int cuda_fail(cudaError_t e, const char* where) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%s)", where, cudaGetErrorName(e),
           cudaGetErrorString(e));
  return MVL_ERR_CUDA;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int elem_size(int dtype) {
  switch (dtype) {
    case MVL_DTYPE_F32: return 4;
    case MVL_DTYPE_F16:
    case MVL_DTYPE_BF16: return 2;
  }
  return 0;
}

int ilog2_exact(long long v) {  // log2(v) if v is a power of two, else -1
  if (v <= 0 || (v & (v - 1))) return -1;
  int s = 0;
  while ((1ll << s) < v) ++s;
  return s;
}

// SM count of the current device, cached per device ordinal.
int current_sm_count(int* out) {
  static int cache[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev >= 0 && dev < 64 && cache[dev] > 0) {
    *out = cache[dev];
    return MVL_OK;
  }
  int n = 0;
  e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
  if (dev >= 0 && dev < 64) cache[dev] = n;
  *out = n;
  return MVL_OK;
}

// Head tile: how many query heads of a GQA group one CTA carries (1, 2, 4 or 8).
int pick_gt(int G) {
  if (G >= 8) return 8;
  if (G > 2) return 4;
  return G;  // 1 or 2
}

int tile_rows(int dtype, int head_dim, int gt) {
  switch (dtype) {
    case MVL_DTYPE_F32: return mvl::decode_tile_rows<float>(head_dim, gt);
    case MVL_DTYPE_F16: return mvl::decode_tile_rows<__half>(head_dim, gt);
    case MVL_DTYPE_BF16: return mvl::decode_tile_rows<__nv_bfloat16>(head_dim, gt);
  }
  return -1;
}

struct Plan {
  int gt, n_gt, tr, chunk, max_splits, bs_shift;
  long long max_ctx;
};

constexpr int kWarps = 4;
constexpr int kMaxChunk = 4096;

// Shape checks shared by the planning entry points and mvl_paged_decode.
int make_plan(long long batch, int H, int KVH, int D, long long max_num_blocks, int block_size,
              int dtype, int chunk_req, int num_sms, Plan* pl, bool mma = false) {
  if (elem_size(dtype) == 0) return MVL_ERR_DTYPE;
  if (batch < 0 || H <= 0 || KVH <= 0 || H % KVH != 0 || max_num_blocks <= 0) return MVL_ERR_SHAPE;
  if (D != 64 && D != 128) return MVL_ERR_SHAPE;
  const int bs_shift = ilog2_exact(block_size);
  if (bs_shift < 0) return MVL_ERR_SHAPE;
  const long long max_ctx = max_num_blocks * (long long)block_size;
  if (max_ctx >= (1ll << 31) || batch > 65535) return MVL_ERR_LIMIT;
  const int G = H / KVH;
  pl->gt = mma ? 8 : pick_gt(G);  // the tensor-core kernel always carries 8 heads (N of the MMA)
  pl->n_gt = (G + pl->gt - 1) / pl->gt;
  pl->tr = mma ? 16 : tile_rows(dtype, D, pl->gt);
  if (pl->tr <= 0) return MVL_ERR_SHAPE;
  pl->bs_shift = bs_shift;
  pl->max_ctx = max_ctx;
  const int cta_rows = pl->tr * kWarps;  // tokens one CTA consumes per round of its warps
  int chunk;
  if (chunk_req > 0) {
    chunk = ((chunk_req + cta_rows - 1) / cta_rows) * cta_rows;
  } else {
    // A CTA pays ~2 us to start (q, table entry, first tile) and the last CTAs to finish leave the
    // chip partly idle for about half a CTA lifetime; the total is smallest when the lifetime is
    // sqrt(2 * 2 us * T), T = the kernel's duration at streaming rate for the worst-case context
    // length.  One SM streams ~44 KB/us (6.5 TB/s / 148) shared by its 3 resident CTAs.
    // Reproduces the measured optima (scripts/sweep_chunks.sh, scripts/sweep_shard.sh): C3 2048,
    // C3 sharded 4- and 8-way 1024, C2 512, C4/C5 4096.
    const double units = (double)(batch > 0 ? batch : 1) * KVH * pl->n_gt;
    const double tok_bytes = 2.0 * D * elem_size(dtype);
    const double rate_sm = 44000.0;  // bytes per microsecond per SM
    const double t_est_us = units * (double)max_ctx * tok_bytes / (rate_sm * num_sms);
    const double life_us = sqrt(4.0 * t_est_us);
    // The tensor-core kernel streams a chunk in less time, so its fixed costs weigh more: its
    // measured optima (scripts/sweep_small.py, scripts/sweep_mma.sh) sit one step higher.
    const double want = life_us * rate_sm / (3.0 * tok_bytes) * (mma ? 2.0 : 1.0);
    chunk = cta_rows;
    while (chunk * 2 <= kMaxChunk && (double)chunk * 1.41421356 < want) chunk *= 2;
    // Wave quantisation (scripts/sweep_small.py): a grid of one to two waves of resident CTAs runs
    // a nearly empty second wave; if half as many, twice as long CTAs still cover half a wave, take
    // those (B=1, ctx 16k: 25.9 -> 21.2 us; B=4, ctx 4k: 21.6 -> 19.0 us).
    const double slots = (mma ? 4.0 : 3.0) * num_sms;  // resident CTAs
    auto ctas = [&](long long c) { return units * (double)((max_ctx + c - 1) / c); };
    if (chunk * 2 <= kMaxChunk && ctas(chunk) <= 2.0 * slots && ctas(chunk * 2) >= 0.5 * slots) chunk *= 2;
  }
  if ((long long)chunk > ((max_ctx + cta_rows - 1) / cta_rows) * cta_rows)
    chunk = (int)(((max_ctx + cta_rows - 1) / cta_rows) * cta_rows);
  pl->chunk = chunk;
  const long long splits = (max_ctx + chunk - 1) / chunk;
  if (splits > 65535) return MVL_ERR_LIMIT;
  pl->max_splits = (int)splits;
  return MVL_OK;
}

// Workspace layout: [arrival counters, fixed kCounterBytes][share meta, 16 B x batch rounded to
// 256][part_o fp32][part_ml float2].
// The counter region has a fixed size and position so that it never overlaps another call's
// partials, whatever the shapes: it only ever holds counters, which every call leaves at zero.
constexpr size_t kCounterBytes = 256u << 10;  // 65536 counters

// leader | gsize | gshare | lcpt | work-list leaders | work-list chunk offsets (batch ints each) + 2 counts
size_t share_meta_bytes(long long batch) { return (((size_t)batch * 24 + 8 + 255) / 256) * 256; }

size_t plan_workspace(const Plan& pl, long long batch, int H, int D) {
  if (pl.max_splits <= 1) return 0;
  const size_t rows = (size_t)batch * H * pl.max_splits;
  return kCounterBytes + share_meta_bytes(batch) + rows * D * sizeof(float) + rows * sizeof(float2);
}

// The persistent stream kernel (paged_decode_stream.cuh): resident CTAs and workspace.  It is a
// candidate only when the caller left the split size to the library (chunk_tokens == 0).
constexpr int kStreamCtasPerSm = 4;
bool stream_eligible(long long batch, int KVH, int n_gt, long long max_ctx, int dtype, int chunk_req) {
  if (chunk_req != 0 || elem_size(dtype) != 2 || batch < 1 || batch > mvl::kStreamBatchLimit) return false;
  const long long units = batch * KVH * n_gt;
  if (units * (long long)sizeof(int) > (long long)kCounterBytes) return false;
  return units * ((max_ctx + 63) / 64) < (1ll << 31);  // stream positions fit comfortably
}
size_t stream_workspace(long long batch, int D, int num_sms) {
  return kCounterBytes + share_meta_bytes(batch) + mvl::stream_workspace_partials(kStreamCtasPerSm * num_sms, D);
}

// ---- tensor maps for the TMA load path: [num_slots][KVH][D] with a [tile_rows][1][D] box
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {  // from the driver, no link-time dependency on libcuda
  static EncodeTiledFn fn = [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      ptr = nullptr;
    return reinterpret_cast<EncodeTiledFn>(ptr);
  }();
  return fn;
}

bool make_cache_map(CUtensorMap* map, const void* cache, int dtype, int D, int KVH, long long num_slots,
                    int tile_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return false;
  const CUtensorMapDataType dt = dtype == MVL_DTYPE_F32    ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : dtype == MVL_DTYPE_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                          : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const cuuint64_t es = elem_size(dtype);
  const cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)KVH, (cuuint64_t)num_slots};
  const cuuint64_t strides[2] = {(cuuint64_t)D * es, (cuuint64_t)KVH * D * es};  // bytes, dims 1..2
  const cuuint32_t box[3] = {(cuuint32_t)D, 1u, (cuuint32_t)tile_rows};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  return enc(map, dt, 3, const_cast<void*>(cache), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Streams and events of mvl_decode_step_host_overlapped, created once per device (no device memory).
constexpr int kMaxPieces = 8;
struct HostPipe {
  cudaStream_t st[kMaxPieces];
  cudaEvent_t start, done[kMaxPieces];
  bool ready;
};
std::mutex g_pipe_mutex;
HostPipe g_pipes[64] = {};

int pipe_for_current_device(HostPipe** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= 64) return MVL_ERR_LIMIT;
  std::lock_guard<std::mutex> lock(g_pipe_mutex);
  HostPipe& hp = g_pipes[dev];
  if (!hp.ready) {
    for (int i = 0; i < kMaxPieces; ++i) {
      e = cudaStreamCreateWithFlags(&hp.st[i], cudaStreamNonBlocking);
      if (e != cudaSuccess) return cuda_fail(e, "cudaStreamCreateWithFlags");
      e = cudaEventCreateWithFlags(&hp.done[i], cudaEventDisableTiming);
      if (e != cudaSuccess) return cuda_fail(e, "cudaEventCreateWithFlags");
    }
    e = cudaEventCreateWithFlags(&hp.start, cudaEventDisableTiming);
    if (e != cudaSuccess) return cuda_fail(e, "cudaEventCreateWithFlags");
    hp.ready = true;
  }
  *out = &hp;
  return MVL_OK;
}

unsigned long long* g_dbg_times = nullptr;  // mvl_debug_set_times

struct PeerSync {  // completion signal of the peer-store gather (peer_sync.cu)
  void* const* blocks;  // every rank's control block as mapped on this device
  int my_rank;
};

struct StoreArgs {  // the step's new K/V rows for the fused store (mvl_decode_step)
  const void* key;
  const void* value;
  const void* slot_mapping;
  int slot_is_i64;
  long long key_token_stride, value_token_stride;  // elements
};

}  // namespace

namespace mvl {
// Tensor map over a packed (tokens, heads, D) 16-bit tensor (prefill.h).
int make_packed_map(CUtensorMap* map, const void* base, int dtype, int D, int heads, long long tokens,
                    long long token_stride, int box_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return -1;
  const CUtensorMapDataType dt =
      dtype == MVL_DTYPE_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)heads, (cuuint64_t)tokens};
  const cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)token_stride * 2};  // bytes, dims 1..2
  const cuuint32_t box[3] = {64u, 1u, (cuuint32_t)box_rows};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  return (int)enc(map, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}
// Tensor map over q (batch, heads, D) whose box is one sub-group's rows for one kv head: 64 elements x
// `box_heads` consecutive heads x `box_seqs` consecutive sequences (row = sequence-major, head-minor).
int make_group_q_map(CUtensorMap* map, const void* q, int dtype, int D, int heads, long long batch, int box_heads,
                     int box_seqs) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return -1;
  const CUtensorMapDataType dt =
      dtype == MVL_DTYPE_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)heads, (cuuint64_t)batch};
  const cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)heads * D * 2};
  const cuuint32_t box[3] = {64u, (cuuint32_t)box_heads, (cuuint32_t)box_seqs};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  return (int)enc(map, dt, 3, const_cast<void*>(q), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}
int fail_text(const char* where, int code) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: code %d", where, code);
  return MVL_ERR_CUDA;
}
int fail_cuda(int cuda_error, const char* where) { return cuda_fail((cudaError_t)cuda_error, where); }
}  // namespace mvl

extern "C" {

int mvl_abi_version(void) { return MVL_ABI_VERSION; }

const char* mvl_strerror(int status) {
  switch (status) {
    case MVL_OK: return "ok";
    case MVL_ERR_NULL_POINTER: return "a required pointer is NULL";
    case MVL_ERR_DTYPE: return "unsupported element dtype (expect f32, f16 or bf16)";
    case MVL_ERR_SHAPE:
      return "unsupported shape (head_dim in {64,128}, num_heads % num_kv_heads == 0, "
             "power-of-two block_size, positive sizes)";
    case MVL_ERR_ALIGNMENT: return "pointer or stride not 16-byte aligned";
    case MVL_ERR_WORKSPACE: return "workspace is NULL or smaller than mvl_paged_decode_workspace_bytes()";
    case MVL_ERR_CUDA: return "CUDA error (see mvl_last_cuda_error())";
    case MVL_ERR_LIMIT: return "size exceeds a fixed limit (batch <= 65535, context < 2^31, splits <= 65535)";
  }
  return "unknown mvl_status";
}

const char* mvl_last_cuda_error(void) { return g_cuda_err; }

int mvl_store_kvcache(const void* key, const void* value, void* k_cache, void* v_cache,
                      const void* slot_mapping, int slot_is_i64, int64_t num_tokens,
                      int num_kv_heads, int head_dim, int64_t key_token_stride,
                      int64_t value_token_stride, int64_t num_slots, int elem_dtype,
                      void* stream) {
  const int es = elem_size(elem_dtype);
  if (es == 0) return MVL_ERR_DTYPE;
  if (num_tokens < 0 || num_kv_heads <= 0 || head_dim <= 0 || num_slots < 0) return MVL_ERR_SHAPE;
  if (num_tokens == 0) return MVL_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping) return MVL_ERR_NULL_POINTER;
  const long long row_bytes = (long long)num_kv_heads * head_dim * es;
  if (row_bytes % 16 != 0 || row_bytes / 16 > 0x7fffffffll) return MVL_ERR_SHAPE;
  if (key_token_stride < (long long)num_kv_heads * head_dim ||
      value_token_stride < (long long)num_kv_heads * head_dim)
    return MVL_ERR_SHAPE;
  if ((key_token_stride * es) % 16 != 0 || (value_token_stride * es) % 16 != 0 ||
      !aligned16(key) || !aligned16(value) || !aligned16(k_cache) || !aligned16(v_cache))
    return MVL_ERR_ALIGNMENT;
  int sms = 0;
  int st = current_sm_count(&sms);
  if (st != MVL_OK) return st;
  mvl::launch_store_kvcache(key, value, k_cache, v_cache, slot_mapping, slot_is_i64, num_tokens,
                            (int)(row_bytes / 16), key_token_stride * es / 16,
                            value_token_stride * es / 16, num_slots, sms, (cudaStream_t)stream);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "store_kvcache launch");
  return MVL_OK;
}

int mvl_paged_decode_default_chunk(int64_t batch, int num_heads, int num_kv_heads, int head_dim,
                                   int64_t max_num_blocks, int block_size, int elem_dtype) {
  int sms = 0;
  int st = current_sm_count(&sms);
  if (st != MVL_OK) return st;
  Plan pl;  // the plan of the default kernel: tensor cores for 16-bit caches
  st = make_plan(batch, num_heads, num_kv_heads, head_dim, max_num_blocks, block_size, elem_dtype, 0,
                 sms, &pl, elem_size(elem_dtype) == 2);
  return st != MVL_OK ? st : pl.chunk;
}

size_t mvl_paged_decode_workspace_bytes(int64_t batch, int num_heads, int num_kv_heads,
                                        int head_dim, int64_t max_num_blocks, int block_size,
                                        int elem_dtype, int chunk_tokens) {
  int sms = 0;
  if (current_sm_count(&sms) != MVL_OK) return 0;
  Plan pl;
  if (make_plan(batch, num_heads, num_kv_heads, head_dim, max_num_blocks, block_size, elem_dtype,
                chunk_tokens, sms, &pl) != MVL_OK)
    return 0;
  size_t need = plan_workspace(pl, batch, num_heads, head_dim);
  if (elem_size(elem_dtype) == 2 &&  // the tensor-core path may split differently: cover both
      make_plan(batch, num_heads, num_kv_heads, head_dim, max_num_blocks, block_size, elem_dtype,
                chunk_tokens, sms, &pl, true) == MVL_OK) {
    const size_t alt = plan_workspace(pl, batch, num_heads, head_dim);
    if (alt > need) need = alt;
    if (stream_eligible(batch, num_kv_heads, pl.n_gt, pl.max_ctx, elem_dtype, chunk_tokens)) {
      const size_t st = stream_workspace(batch, head_dim, sms);
      if (st > need) need = st;
    }
  }
  return need;
}

static int decode_impl(const StoreArgs* st_args, void* const* outs, int n_out, int h_out, int head_off,
                       const void* query,
                       const void* k_cache, const void* v_cache,
                     const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                     int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                     int num_heads, int num_kv_heads, int head_dim, int block_size,
                     int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                     size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream,
                     const PeerSync* sync = nullptr);

static int decode_impl(const StoreArgs* st_args, void* const* outs, int n_out, int h_out, int head_off,
                       const void* query,
                       const void* k_cache, const void* v_cache,
                     const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                     int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                     int num_heads, int num_kv_heads, int head_dim, int block_size,
                     int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                     size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream,
                     const PeerSync* sync) {
  int sms = 0;
  int st = current_sm_count(&sms);
  if (st != MVL_OK) return st;
  const bool mma = !(flags & MVL_DECODE_CUDA_CORES) && !(flags & MVL_DECODE_LOAD_TMA) &&
                   elem_size(elem_dtype) == 2;
  Plan pl;
  st = make_plan(batch, num_heads, num_kv_heads, head_dim, max_num_blocks, block_size, elem_dtype,
                 chunk_tokens, sms, &pl, mma);
  if (st != MVL_OK) return st;
  if (sync) {
    if (!sync->blocks || sync->my_rank < 0 || sync->my_rank >= n_out) return MVL_ERR_SHAPE;
    for (int r = 0; r < n_out && r < 8; ++r)
      if (!sync->blocks[r]) return MVL_ERR_NULL_POINTER;
    if (flags & (MVL_DECODE_SKIP_COMBINE | MVL_DECODE_ONLY_COMBINE)) return MVL_ERR_SHAPE;
  }
  cudaStream_t s = (cudaStream_t)stream;
  // with a completion signal even an empty step must announce itself
  auto signal_after = [&]() -> int {
    if (!sync) return MVL_OK;
    unsigned* blocks[8] = {};
    for (int r = 0; r < n_out; ++r) blocks[r] = static_cast<unsigned*>(sync->blocks[r]);
    const int e = mvl::launch_peer_signal(blocks[sync->my_rank], blocks, n_out, sync->my_rank, s);
    return e == 0 ? MVL_OK : cuda_fail((cudaError_t)e, "peer_signal launch");
  };
  if (batch == 0) return signal_after();
  if (num_blocks <= 0 || block_table_stride < max_num_blocks) return MVL_ERR_SHAPE;
  if (!outs || !query || !k_cache || !v_cache || !block_tables || !context_lens)
    return MVL_ERR_NULL_POINTER;
  if (n_out < 1 || n_out > 8) return MVL_ERR_LIMIT;
  if (head_off < 0 || head_off + num_heads > h_out) return MVL_ERR_SHAPE;
  for (int r = 0; r < n_out; ++r) {
    if (!outs[r]) return MVL_ERR_NULL_POINTER;
    if (!aligned16(outs[r])) return MVL_ERR_ALIGNMENT;
  }
  if (!aligned16(query) || !aligned16(k_cache) || !aligned16(v_cache) ||
      (reinterpret_cast<uintptr_t>(block_tables) & 3u) ||
      (reinterpret_cast<uintptr_t>(context_lens) & (ctx_is_i64 ? 7u : 3u)))
    return MVL_ERR_ALIGNMENT;
  const size_t need = plan_workspace(pl, batch, num_heads, head_dim);
  if (need > 0 && (!workspace || workspace_bytes < need)) return MVL_ERR_WORKSPACE;
  if (need > 0 && !aligned16(workspace)) return MVL_ERR_ALIGNMENT;
  // Persistent stream kernel instead of the chunk grid?  Only where its conditions hold and the
  // workspace the caller sized with mvl_paged_decode_workspace_bytes() is there.
  bool use_stream = false;
  if (mma && (flags & MVL_DECODE_FUSED_COMBINE) &&
      !(flags & (MVL_DECODE_SPLIT_GRID | MVL_DECODE_SHARE_PREFIX | MVL_DECODE_SKIP_COMBINE |
                 MVL_DECODE_ONLY_COMBINE)) &&
      stream_eligible(batch, num_kv_heads, pl.n_gt, pl.max_ctx, elem_dtype, chunk_tokens) && workspace &&
      aligned16(workspace) && workspace_bytes >= stream_workspace(batch, head_dim, sms)) {
    // Measured on B200 (profiles/r02_sweep_stream_vs_grid.jsonl, r02_sweep_small.jsonl): the chunk grid is
    // as fast or faster on every shape but the one-kv-head shard of a large batch (1-2 % there): CTAs with
    // equal work finish +-10 % apart, and the hardware's CTA scheduler, handing out many small CTAs as slots
    // free up, evens that out where fixed pieces cannot.  So the stream kernel runs only when asked for.
    use_stream = (flags & MVL_DECODE_STREAM) != 0;
  }

  mvl::DecodeParams p;
  memset(&p, 0, sizeof(p));
  for (int r = 0; r < n_out; ++r) p.outs[r] = outs[r];
  p.n_out = n_out;
  p.H_out = h_out;
  p.head_off = head_off;
  p.q = query;
  p.kc = k_cache;
  p.vc = v_cache;
  p.bt = block_tables;
  p.ctx = context_lens;
  const size_t rows = (size_t)batch * num_heads * pl.max_splits;
  if (use_stream) {
    char* base = reinterpret_cast<char*>(workspace);
    const size_t head_bytes = kCounterBytes + share_meta_bytes(batch);
    const size_t slots = (size_t)kStreamCtasPerSm * sms * 2;
    p.part_o = reinterpret_cast<float*>(base + head_bytes);
    p.part_ml = reinterpret_cast<float2*>(base + head_bytes + slots * 8 * head_dim * sizeof(float));
    p.counters = reinterpret_cast<int*>(base);
  } else if (need) {
    char* base = reinterpret_cast<char*>(workspace);
    const size_t head_bytes = kCounterBytes + share_meta_bytes(batch);
    p.part_o = reinterpret_cast<float*>(base + head_bytes);
    p.part_ml = reinterpret_cast<float2*>(base + head_bytes + rows * head_dim * sizeof(float));
    const size_t n_counters = (size_t)batch * num_kv_heads * pl.n_gt;
    if ((flags & MVL_DECODE_FUSED_COMBINE) && n_counters * sizeof(int) <= kCounterBytes &&
        !(flags & (MVL_DECODE_SKIP_COMBINE | MVL_DECODE_ONLY_COMBINE)))
      p.counters = reinterpret_cast<int*>(base);
  }
  if (st_args) {
    const int es = elem_size(elem_dtype);
    if (!st_args->key || !st_args->value || !st_args->slot_mapping) return MVL_ERR_NULL_POINTER;
    const long long dense = (long long)num_kv_heads * head_dim;
    if (st_args->key_token_stride < dense || st_args->value_token_stride < dense) return MVL_ERR_SHAPE;
    if ((st_args->key_token_stride * es) % 16 || (st_args->value_token_stride * es) % 16 ||
        !aligned16(st_args->key) || !aligned16(st_args->value))
      return MVL_ERR_ALIGNMENT;
    p.k_new = st_args->key;
    p.v_new = st_args->value;
    p.slot_map = st_args->slot_mapping;
    p.slot_is_i64 = st_args->slot_is_i64 ? 1 : 0;
    p.k_new_stride = st_args->key_token_stride * es;
    p.v_new_stride = st_args->value_token_stride * es;
    p.num_slots = num_blocks * (long long)block_size;
  }
  p.bt_stride = block_table_stride;
  p.max_ctx = pl.max_ctx;
  p.ctx_is_i64 = ctx_is_i64 ? 1 : 0;
  p.H = num_heads;
  p.KVH = num_kv_heads;
  p.G = num_heads / num_kv_heads;
  p.n_gt = pl.n_gt;
  p.bs_shift = pl.bs_shift;
  p.chunk = pl.chunk;
  p.max_splits = pl.max_splits;
  p.qscale = scale * 1.4426950408889634f;
  p.gp_shift = 3;  // 8 columns = 8 heads of one sequence unless prefix sharing packs several
  p.mm_narrow = 1;
  p.wide = 0;
  p.overlap_prev = (flags & MVL_DECODE_OVERLAP_PREV) ? 1 : 0;
  p.dbg_times = g_dbg_times;
  // Prefix sharing: tensor-core kernel, fewer than 8 heads per KV head (spare columns), more than
  // one sequence, and a workspace to hold the grouping.  Two tiny launches find the groups.
  if ((flags & MVL_DECODE_SHARE_PREFIX) && mma && need && batch > 1 && p.G <= 4 &&
      !(flags & MVL_DECODE_ONLY_COMBINE)) {
    int gp = 1;
    while (gp < p.G) gp <<= 1;
    p.gp_shift = gp == 1 ? 0 : (gp == 2 ? 1 : 2);
    p.mm_narrow = 8 / gp;
    // wide form: sub-groups of up to 4 x (8 / gp) sequences, one CTA per shared chunk reads its tiles
    // once for all of them (paged_decode_group.cuh); needs whole 16-token tiles inside a page
    p.wide = (!(flags & MVL_DECODE_SHARE_NARROW) && block_size % 16 == 0 && pl.chunk % 64 == 0) ? 1 : 0;
    int* meta = reinterpret_cast<int*>(reinterpret_cast<char*>(workspace) + kCounterBytes);
    const int e = mvl::launch_prefix_share(block_tables, block_table_stride, context_lens, p.ctx_is_i64,
                                           batch, pl.max_ctx, pl.bs_shift, pl.chunk,
                                           p.wide ? 4 * p.mm_narrow : p.mm_narrow, p.mm_narrow, meta, s);
    if (e != 0) return cuda_fail((cudaError_t)e, "prefix_share launch");
    p.share_meta = meta;
  }

  bool signalled_in_kernel = false;
  auto set_sync = [&]() {
    for (int r = 0; r < n_out; ++r) p.sync_peers[r] = static_cast<unsigned*>(sync->blocks[r]);
    p.sync_local = p.sync_peers[sync->my_rank];
    p.my_rank = sync->my_rank;
  };
  if (use_stream) {
    // no more CTAs than 64-token rows the launch can possibly hold
    long long ctas = (long long)kStreamCtasPerSm * sms;
    const long long max_rows = batch * num_kv_heads * pl.n_gt * ((pl.max_ctx + 63) / 64);
    if (ctas > max_rows) ctas = max_rows;
    if (sync) set_sync();  // signalled from the kernel's own epilogue
    const int e = elem_dtype == MVL_DTYPE_F16
                      ? mvl::launch_decode_stream_mma<__half>(p, head_dim, (int)batch, (int)ctas, s)
                      : mvl::launch_decode_stream_mma<__nv_bfloat16>(p, head_dim, (int)batch, (int)ctas, s);
    if (e == -1000) return MVL_ERR_SHAPE;
    if (e != 0) return cuda_fail((cudaError_t)e, "paged_decode_stream launch");
    return MVL_OK;
  }
  if (!(flags & MVL_DECODE_ONLY_COMBINE)) {
    dim3 grid((unsigned)(num_kv_heads * pl.n_gt), (unsigned)batch, (unsigned)pl.max_splits);
    CUtensorMap maps[2];
    const CUtensorMap* use_maps = nullptr;
    if ((flags & MVL_DECODE_LOAD_TMA) && block_size % pl.tr == 0) {
      const long long slots = num_blocks * (long long)block_size;
      if (make_cache_map(&maps[0], k_cache, elem_dtype, head_dim, num_kv_heads, slots, pl.tr) &&
          make_cache_map(&maps[1], v_cache, elem_dtype, head_dim, num_kv_heads, slots, pl.tr))
        use_maps = maps;
    }
    int e = -1000;
    // Prefix sharing, wide form: the shared chunks of large sub-groups go to the tcgen05 group kernel (one
    // CTA per SM, persistent) where its tiling holds — 128-token tiles inside chunks, whole 16-token pieces
    // inside pages — else to the mma.sync one.  The tcgen05 kernel is launched FIRST and the split kernel
    // as its programmatic dependent: its big uniform CTAs start on every SM at once, and the split kernel's
    // many small ones fill each SM as a group CTA leaves it (launched behind the split kernel it started
    // SM by SM over 80 us and ended a third round nearly alone: 0.396 ms on C5 instead of what is measured now).
    const bool group_wide = mma && p.share_meta != nullptr && p.wide;
    const bool group_tc5 = group_wide && !(flags & MVL_DECODE_GROUP_MMA) && pl.chunk % 128 == 0 &&
                           block_size % 16 == 0 && num_blocks * (long long)block_size < (1ll << 31);
    if (group_tc5) {
      e = mvl::launch_decode_group_tc5(p, head_dim, batch, block_size, num_blocks, sms, elem_dtype, s);
      if (e == -1000) return MVL_ERR_SHAPE;
      if (e != 0) return e;  // text recorded by the launcher
      p.beside_group = 1;
    }
    // the tensor-core split kernel announces the step itself when no later kernel writes result rows
    signalled_in_kernel = sync && mma && p.share_meta == nullptr && !(flags & MVL_DECODE_SKIP_COMBINE) &&
                          (pl.max_splits <= 1 || p.counters != nullptr);
    if (signalled_in_kernel) set_sync();
    if (mma) {
      e = elem_dtype == MVL_DTYPE_F16 ? mvl::launch_decode_split_mma<__half>(p, head_dim, grid, s)
                                      : mvl::launch_decode_split_mma<__nv_bfloat16>(p, head_dim, grid, s);
    } else
    switch (elem_dtype) {
      case MVL_DTYPE_F32: e = mvl::launch_decode_split<float>(p, use_maps, head_dim, pl.gt, grid, s); break;
      case MVL_DTYPE_F16: e = mvl::launch_decode_split<__half>(p, use_maps, head_dim, pl.gt, grid, s); break;
      case MVL_DTYPE_BF16:
        e = mvl::launch_decode_split<__nv_bfloat16>(p, use_maps, head_dim, pl.gt, grid, s);
        break;
    }
    if (e == -1000) return MVL_ERR_SHAPE;
    if (e != 0) return cuda_fail((cudaError_t)e, "paged_decode_split launch");
    if (group_wide && !group_tc5) {
      e = elem_dtype == MVL_DTYPE_F16 ? mvl::launch_decode_group_mma<__half>(p, head_dim, grid, s)
                                      : mvl::launch_decode_group_mma<__nv_bfloat16>(p, head_dim, grid, s);
      if (e == -1000) return MVL_ERR_SHAPE;
      if (e != 0) return cuda_fail((cudaError_t)e, "paged_decode_group launch");
    }
  }
  if (pl.max_splits > 1 && !(flags & MVL_DECODE_SKIP_COMBINE) && p.counters == nullptr) {
    const long long bh = (long long)batch * num_heads;
    int e = -1000;
    switch (elem_dtype) {
      case MVL_DTYPE_F32: e = mvl::launch_decode_combine<float>(p, head_dim, bh, s); break;
      case MVL_DTYPE_F16: e = mvl::launch_decode_combine<__half>(p, head_dim, bh, s); break;
      case MVL_DTYPE_BF16: e = mvl::launch_decode_combine<__nv_bfloat16>(p, head_dim, bh, s); break;
    }
    if (e == -1000) return MVL_ERR_SHAPE;
    if (e != 0) return cuda_fail((cudaError_t)e, "paged_decode_combine launch");
  }
  return signalled_in_kernel ? MVL_OK : signal_after();  // else: one-warp signal launch behind the kernels
}

int mvl_paged_decode(void* out, const void* query, const void* k_cache, const void* v_cache,
                     const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                     int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                     int num_heads, int num_kv_heads, int head_dim, int block_size,
                     int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                     size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream) {
  void* outs[1] = {out};
  return decode_impl(nullptr, outs, 1, num_heads, 0, query, k_cache, v_cache, block_tables, context_lens,
                     ctx_is_i64, batch, max_num_blocks, block_table_stride, num_heads, num_kv_heads,
                     head_dim, block_size, num_blocks, scale, elem_dtype, workspace, workspace_bytes,
                     chunk_tokens, flags, stream);
}

int mvl_paged_decode_peers(void* const* out_peers, int num_peers, int total_heads, int head_offset,
                           const void* query, const void* k_cache, const void* v_cache,
                           const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                           int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                           int num_heads, int num_kv_heads, int head_dim, int block_size,
                           int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                           size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream) {
  return decode_impl(nullptr, out_peers, num_peers, total_heads, head_offset, query, k_cache, v_cache,
                     block_tables, context_lens, ctx_is_i64, batch, max_num_blocks,
                     block_table_stride, num_heads, num_kv_heads, head_dim, block_size, num_blocks,
                     scale, elem_dtype, workspace, workspace_bytes, chunk_tokens, flags, stream);
}

int mvl_decode_step(void* out, const void* query, const void* key, const void* value, void* k_cache,
                    void* v_cache, const void* slot_mapping, const int32_t* block_tables,
                    const void* context_lens, int idx_is_i64, int64_t batch, int64_t max_num_blocks,
                    int64_t block_table_stride, int64_t key_token_stride, int64_t value_token_stride,
                    int num_heads, int num_kv_heads, int head_dim, int block_size,
                    int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                    size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream) {
  StoreArgs st{key, value, slot_mapping, idx_is_i64, key_token_stride, value_token_stride};
  void* outs[1] = {out};
  return decode_impl(&st, outs, 1, num_heads, 0, query, k_cache, v_cache, block_tables, context_lens,
                     idx_is_i64, batch, max_num_blocks, block_table_stride, num_heads, num_kv_heads,
                     head_dim, block_size, num_blocks, scale, elem_dtype, workspace, workspace_bytes,
                     chunk_tokens, flags, stream);
}

int mvl_decode_step_peers(void* const* out_peers, int num_peers, int total_heads, int head_offset,
                          const void* query, const void* key, const void* value, void* k_cache,
                          void* v_cache, const void* slot_mapping, const int32_t* block_tables,
                          const void* context_lens, int idx_is_i64, int64_t batch,
                          int64_t max_num_blocks, int64_t block_table_stride,
                          int64_t key_token_stride, int64_t value_token_stride, int num_heads,
                          int num_kv_heads, int head_dim, int block_size, int64_t num_blocks,
                          float scale, int elem_dtype, void* workspace, size_t workspace_bytes,
                          int chunk_tokens, unsigned flags, void* stream) {
  StoreArgs st{key, value, slot_mapping, idx_is_i64, key_token_stride, value_token_stride};
  return decode_impl(&st, out_peers, num_peers, total_heads, head_offset, query, k_cache, v_cache,
                     block_tables, context_lens, idx_is_i64, batch, max_num_blocks,
                     block_table_stride, num_heads, num_kv_heads, head_dim, block_size, num_blocks,
                     scale, elem_dtype, workspace, workspace_bytes, chunk_tokens, flags, stream);
}

int mvl_decode_step_peers_sync(void* const* out_peers, void* const* sync_peers, int num_peers, int my_rank,
                               int total_heads, int head_offset, const void* query, const void* key,
                               const void* value, void* k_cache, void* v_cache, const void* slot_mapping,
                               const int32_t* block_tables, const void* context_lens, int idx_is_i64,
                               int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                               int64_t key_token_stride, int64_t value_token_stride, int num_heads,
                               int num_kv_heads, int head_dim, int block_size, int64_t num_blocks,
                               float scale, int elem_dtype, void* workspace, size_t workspace_bytes,
                               int chunk_tokens, unsigned flags, void* stream) {
  if (!sync_peers) return MVL_ERR_NULL_POINTER;
  if (num_peers < 1 || num_peers > 8) return MVL_ERR_LIMIT;
  PeerSync ps{sync_peers, my_rank};
  StoreArgs st{key, value, slot_mapping, idx_is_i64, key_token_stride, value_token_stride};
  const bool with_store = key != nullptr || value != nullptr || slot_mapping != nullptr;
  return decode_impl(with_store ? &st : nullptr, out_peers, num_peers, total_heads, head_offset, query,
                     k_cache, v_cache, block_tables, context_lens, idx_is_i64, batch, max_num_blocks,
                     block_table_stride, num_heads, num_kv_heads, head_dim, block_size, num_blocks, scale,
                     elem_dtype, workspace, workspace_bytes, chunk_tokens, flags, stream, &ps);
}

void mvl_debug_set_times(void* buf) { g_dbg_times = static_cast<unsigned long long*>(buf); }

int mvl_peer_wait(void* sync_local, int num_peers, int64_t timeout_ms, void* stream) {
  if (!sync_local) return MVL_ERR_NULL_POINTER;
  if (num_peers < 1 || num_peers > 8) return MVL_ERR_LIMIT;
  if (timeout_ms <= 0) timeout_ms = 10000;
  const int e = mvl::launch_peer_wait(static_cast<unsigned*>(sync_local), num_peers,
                                      (unsigned long long)timeout_ms * 1000000ull, (cudaStream_t)stream);
  return e == 0 ? MVL_OK : cuda_fail((cudaError_t)e, "peer_wait launch");
}

int mvl_decode_step_host(const void* q_host, const void* k_host, const void* v_host,
                         const void* slot_mapping_host, const void* context_lens_host,
                         const int32_t* block_tables_host, void* out_host, void* q_dev,
                         void* k_dev, void* v_dev, void* slot_mapping_dev, void* context_lens_dev,
                         int32_t* block_tables_dev, void* out_dev, void* k_cache, void* v_cache,
                         int idx_is_i64, int64_t batch, int64_t max_num_blocks, int num_heads,
                         int num_kv_heads, int head_dim, int block_size, int64_t num_blocks,
                         float scale, int elem_dtype, void* workspace, size_t workspace_bytes,
                         int chunk_tokens, unsigned flags, void* stream) {
  const int es = elem_size(elem_dtype);
  if (es == 0) return MVL_ERR_DTYPE;
  if (batch < 0 || num_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0 || max_num_blocks <= 0)
    return MVL_ERR_SHAPE;
  if (batch == 0) return MVL_OK;
  if (!q_host || !k_host || !v_host || !slot_mapping_host || !context_lens_host ||
      !block_tables_host || !out_host || !q_dev || !k_dev || !v_dev || !slot_mapping_dev ||
      !context_lens_dev || !block_tables_dev || !out_dev)
    return MVL_ERR_NULL_POINTER;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t qb = (size_t)batch * num_heads * head_dim * es;
  const size_t kb = (size_t)batch * num_kv_heads * head_dim * es;
  const size_t ib = (size_t)batch * (idx_is_i64 ? 8 : 4);
  const size_t tb = (size_t)batch * max_num_blocks * sizeof(int32_t);
  cudaError_t e;
  // Zero-copy options: a device-side pointer equal to its host-side pointer means "this array stays in
  // (pinned, device-mapped) host memory": the kernels read q and the new token's k / v rows from it and
  // write the result rows into it across PCIe themselves, as they go, and no separate copy is enqueued
  // for it.  Slots, lengths and block tables are always copied (the kernels read them many times).
  const bool q_direct = q_dev == q_host, out_direct = out_dev == out_host;
  const bool kv_direct = k_dev == k_host && v_dev == v_host;  // the new token's rows: read once by the fused store
  // When the caller laid the input arrays out back to back (same order, host and device), what has to
  // travel goes in ONE copy instead of up to six.
  auto after = [](const void* p, size_t n) { return static_cast<const char*>(p) + n; };
  const bool meta_packed =
      after(slot_mapping_host, ib) == context_lens_host && after(context_lens_host, ib) == (const void*)block_tables_host &&
      after(slot_mapping_dev, ib) == context_lens_dev && after(context_lens_dev, ib) == (const void*)block_tables_dev;
  const bool kv_packed = !kv_direct && meta_packed && after(k_host, kb) == v_host && after(v_host, kb) == slot_mapping_host &&
                         after(k_dev, kb) == v_dev && after(v_dev, kb) == slot_mapping_dev;
  const bool packed = !q_direct && kv_packed && after(q_host, qb) == k_host && after(q_dev, qb) == k_dev;
#define MVL_H2D(dst, src, n)                                          \
  e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s);        \
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync H2D " #dst)
  if (packed) {
    MVL_H2D(q_dev, q_host, qb + 2 * kb + 2 * ib + tb);
  } else {
    if (!q_direct) { MVL_H2D(q_dev, q_host, qb); }
    if (kv_packed) {
      MVL_H2D(k_dev, k_host, 2 * kb + 2 * ib + tb);
    } else {
      if (!kv_direct) {
        MVL_H2D(k_dev, k_host, kb);
        MVL_H2D(v_dev, v_host, kb);
      }
      if (meta_packed) {
        MVL_H2D(slot_mapping_dev, slot_mapping_host, 2 * ib + tb);
      } else {
        MVL_H2D(slot_mapping_dev, slot_mapping_host, ib);
        MVL_H2D(context_lens_dev, context_lens_host, ib);
        MVL_H2D(block_tables_dev, block_tables_host, tb);
      }
    }
  }
#undef MVL_H2D
  const int st = mvl_decode_step(out_dev, q_dev, k_dev, v_dev, k_cache, v_cache, slot_mapping_dev,
                                 block_tables_dev, context_lens_dev, idx_is_i64, batch, max_num_blocks,
                                 max_num_blocks, (int64_t)num_kv_heads * head_dim,
                                 (int64_t)num_kv_heads * head_dim, num_heads, num_kv_heads, head_dim,
                                 block_size, num_blocks, scale, elem_dtype, workspace, workspace_bytes,
                                 chunk_tokens, MVL_DECODE_FUSED_COMBINE | flags, stream);
  if (st != MVL_OK) return st;
  if (!out_direct) {
    e = cudaMemcpyAsync(out_host, out_dev, qb, cudaMemcpyDeviceToHost, s);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync D2H out");
  }
  return MVL_OK;
}

size_t mvl_decode_step_host_workspace_bytes(int64_t batch, int num_heads, int num_kv_heads, int head_dim,
                                            int64_t max_num_blocks, int block_size, int elem_dtype,
                                            int chunk_tokens, int pieces) {
  if (pieces < 1) pieces = 1;
  if (pieces > kMaxPieces) pieces = kMaxPieces;
  const int64_t per = (batch + pieces - 1) / pieces;
  const size_t one = mvl_paged_decode_workspace_bytes(per, num_heads, num_kv_heads, head_dim, max_num_blocks,
                                                      block_size, elem_dtype, chunk_tokens);
  return ((one + 255) / 256 * 256) * (size_t)pieces;
}

int mvl_decode_step_host_overlapped(const void* q_host, const void* k_host, const void* v_host,
                                    const void* slot_mapping_host, const void* context_lens_host,
                                    const int32_t* block_tables_host, void* out_host, void* q_dev,
                                    void* k_dev, void* v_dev, void* slot_mapping_dev,
                                    void* context_lens_dev, int32_t* block_tables_dev, void* out_dev,
                                    void* k_cache, void* v_cache, int idx_is_i64, int64_t batch,
                                    int64_t max_num_blocks, int num_heads, int num_kv_heads,
                                    int head_dim, int block_size, int64_t num_blocks, float scale,
                                    int elem_dtype, void* workspace, size_t workspace_bytes,
                                    int chunk_tokens, unsigned flags, int pieces, void* stream) {
  const int es = elem_size(elem_dtype);
  if (es == 0) return MVL_ERR_DTYPE;
  if (batch < 0 || num_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0 || max_num_blocks <= 0)
    return MVL_ERR_SHAPE;
  if (pieces < 1 || pieces > kMaxPieces) return MVL_ERR_LIMIT;
  if (batch == 0) return MVL_OK;
  if (!q_host || !k_host || !v_host || !slot_mapping_host || !context_lens_host ||
      !block_tables_host || !out_host || !q_dev || !k_dev || !v_dev || !slot_mapping_dev ||
      !context_lens_dev || !block_tables_dev || !out_dev)
    return MVL_ERR_NULL_POINTER;
  if (pieces > batch) pieces = (int)batch;
  const size_t need = mvl_decode_step_host_workspace_bytes(batch, num_heads, num_kv_heads, head_dim,
                                                           max_num_blocks, block_size, elem_dtype,
                                                           chunk_tokens, pieces);
  if (need > 0 && (!workspace || workspace_bytes < need)) return MVL_ERR_WORKSPACE;
  const size_t region = need / (size_t)pieces;
  HostPipe* hp = nullptr;
  int st = pipe_for_current_device(&hp);
  if (st != MVL_OK) return st;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaEventRecord(hp->start, s);
  if (e != cudaSuccess) return cuda_fail(e, "cudaEventRecord");
  const size_t qrow = (size_t)num_heads * head_dim * es, krow = (size_t)num_kv_heads * head_dim * es;
  const size_t irow = idx_is_i64 ? 8 : 4, trow = (size_t)max_num_blocks * sizeof(int32_t);
  const int64_t per = (batch + pieces - 1) / pieces;
  auto at = [](const void* p, size_t off) { return static_cast<const char*>(p) + off; };
  auto atw = [](void* p, size_t off) { return static_cast<char*>(p) + off; };
#define MVL_H2D_PIECE(dst, src, row)                                                                \
  e = cudaMemcpyAsync(atw(dst, b0 * (row)), at(src, b0 * (row)), nb * (row), cudaMemcpyHostToDevice, \
                      ps);                                                                          \
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync H2D " #dst)
  for (int i = 0; i < pieces; ++i) {
    const size_t b0 = (size_t)i * per;
    if ((int64_t)b0 >= batch) break;
    const size_t nb = (size_t)((int64_t)(b0 + per) <= batch ? per : batch - (int64_t)b0);
    cudaStream_t ps = hp->st[i];
    e = cudaStreamWaitEvent(ps, hp->start, 0);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamWaitEvent");
    MVL_H2D_PIECE(q_dev, q_host, qrow);
    MVL_H2D_PIECE(k_dev, k_host, krow);
    MVL_H2D_PIECE(v_dev, v_host, krow);
    MVL_H2D_PIECE(slot_mapping_dev, slot_mapping_host, irow);
    MVL_H2D_PIECE(context_lens_dev, context_lens_host, irow);
    MVL_H2D_PIECE(block_tables_dev, block_tables_host, trow);
    st = mvl_decode_step(atw(out_dev, b0 * qrow), at(q_dev, b0 * qrow), at(k_dev, b0 * krow),
                         at(v_dev, b0 * krow), k_cache, v_cache, at(slot_mapping_dev, b0 * irow),
                         reinterpret_cast<const int32_t*>(at(block_tables_dev, b0 * trow)),
                         at(context_lens_dev, b0 * irow), idx_is_i64, (int64_t)nb, max_num_blocks,
                         max_num_blocks, (int64_t)num_kv_heads * head_dim,
                         (int64_t)num_kv_heads * head_dim, num_heads, num_kv_heads, head_dim, block_size,
                         num_blocks, scale, elem_dtype, region ? atw(workspace, (size_t)i * region) : nullptr,
                         region, chunk_tokens, MVL_DECODE_FUSED_COMBINE | flags, ps);
    if (st != MVL_OK) return st;
    e = cudaMemcpyAsync(atw(out_host, b0 * qrow), at(out_dev, b0 * qrow), nb * qrow, cudaMemcpyDeviceToHost, ps);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync D2H out");
    e = cudaEventRecord(hp->done[i], ps);
    if (e != cudaSuccess) return cuda_fail(e, "cudaEventRecord");
    e = cudaStreamWaitEvent(s, hp->done[i], 0);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamWaitEvent");
  }
#undef MVL_H2D_PIECE
  return MVL_OK;
}

}  // extern "C"
