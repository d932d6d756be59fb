/*
 * mvl_abi.h — C ABI of libmvl_b200.so: paged-KV decode attention for MinivLLM on B200 (sm_100a).
 *
 * This is the drop-in boundary. The reference has no native layer: its Python host code hands
 * torch tensors to Triton JIT kernels (reference src/myvllm/layers/attention.py:97-108 and
 * :455-471).  Here the same Python host code (minivllm_b200/layers/attention.py) hands raw device
 * pointers and sizes to the functions below through ctypes.  No torch types, no C++ types, no
 * exceptions cross this boundary.
 *
 * Conventions
 *   - Every function returns MVL_OK (0) or a negative mvl_status; nothing throws or aborts.
 *   - The library never allocates, frees or retains device memory.  Scratch memory ("workspace")
 *     is owned by the caller and sized with mvl_paged_decode_workspace_bytes().  The one exception
 *     is mvl_paged_decode_static (end of this file), whose scratch is a fixed array in the
 *     library's own device image.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All work is
 *     enqueued on it; no call synchronises the device or reads device memory on the host, and the
 *     grid depends only on host-known sizes, so calls may be captured in a CUDA graph
 *     (reference src/myvllm/engine/model_runner.py:317-369 captures the decode forward).
 *   - All data pointers must be 16-byte aligned; rows (head_dim * elem size) are multiples of 16 B.
 *   - KV-cache layout is the reference's: (num_blocks, block_size, num_kv_heads, head_dim),
 *     contiguous (reference attention.py:22,52-56; model_runner.py:162).
 */
#ifndef MVL_ABI_H_
#define MVL_ABI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MVL_ABI_VERSION 1

typedef enum mvl_status {
  MVL_OK = 0,
  MVL_ERR_NULL_POINTER = -1,  /* a required pointer is NULL */
  MVL_ERR_DTYPE = -2,         /* elem_dtype is not one of mvl_dtype */
  MVL_ERR_SHAPE = -3,         /* head_dim / head counts / block_size / sizes not supported */
  MVL_ERR_ALIGNMENT = -4,     /* a pointer or stride is not 16-byte aligned */
  MVL_ERR_WORKSPACE = -5,     /* workspace missing or smaller than ..._workspace_bytes() */
  MVL_ERR_CUDA = -6,          /* a CUDA runtime call / launch failed; see mvl_last_cuda_error() */
  MVL_ERR_LIMIT = -7          /* a size exceeds a documented fixed limit */
} mvl_status;

typedef enum mvl_dtype {
  MVL_DTYPE_F32 = 0,
  MVL_DTYPE_F16 = 1,
  MVL_DTYPE_BF16 = 2
} mvl_dtype;

/* Flags for mvl_paged_decode (bit mask). */
#define MVL_DECODE_SKIP_COMBINE 1u /* profiling only: launch the split kernel, not the combine */
#define MVL_DECODE_ONLY_COMBINE 2u /* profiling only: launch the combine kernel alone */
#define MVL_DECODE_FUSED_COMBINE 4u /* merge the chunks inside the split kernel (last CTA to arrive
                                       per sequence and head tile does it): no combine launch.
                                       Needs the first 256 KiB of the workspace to be zero on first
                                       use; every call leaves them zero.  Ignored (separate combine
                                       kernel) when batch * kv-head tiles > 65536. */

#define MVL_DECODE_LOAD_TMA 8u /* fetch K/V tiles with tensor-map TMA (cp.async.bulk.tensor, one
                                  instruction per tile, mbarrier completion) instead of cp.async.
                                  Used when block_size is a multiple of the 16-row (bf16/fp16,
                                  head_dim 128) warp tile; otherwise the cp.async path runs. */

#define MVL_DECODE_CUDA_CORES 16u /* 16-bit caches run their score and P.V products on the tensor
                                     cores by default (mma.sync m16n8k16, fp32 accumulate, p carried
                                     as a 16-bit value plus a 16-bit correction; see DESIGN.md for why
                                     this departs from "no tensor cores").  This flag selects the
                                     packed-fp32 (FFMA2) CUDA-core kernel instead; fp32 caches always
                                     use it. */

#define MVL_DECODE_SHARE_PREFIX 32u /* sequences whose block tables start with the same entries
                                       (prefix caching, reference block_manager.py:67-97) are found
                                       by a device pre-pass (two small launches) and scored together
                                       against ONE copy of the shared K/V pages: with G <= 4 query
                                       heads per KV head the tensor-core tile's 8 columns hold 8/G
                                       sequences (bit-identical results in that form).  Needs the tensor-core
                                       kernel, G <= 4 and a workspace; ignored otherwise.
                                       Sub-groups hold up to 4 x 8/G sequences: chunks shared by more
                                       than 8/G of them go to a second kernel whose four warps hold
                                       different sequences and read each K/V tile once for all;
                                       there results agree with the unshared run to rounding. */
#define MVL_DECODE_SHARE_NARROW 64u /* with SHARE_PREFIX: sub-groups of 8/G sequences only, no second
                                       kernel (measurement switch) */
#define MVL_DECODE_GROUP_MMA 1024u /* with SHARE_PREFIX (wide form): keep the shared chunks of large
                                       sub-groups on the mma.sync group kernel instead of the tcgen05 one
                                       (measurement switch) */
#define MVL_DECODE_SPLIT_GRID 128u /* reserved (the chunk grid is the default); accepted and ignored */
#define MVL_DECODE_STREAM 256u     /* 16-bit caches, chunk_tokens == 0, FUSED_COMBINE set, batch <= 1024:
                                       run the persistent stream kernel — the launch's whole token stream
                                       cut into equal pieces, one per resident CTA (4 x SMs) — instead of
                                       one CTA per (sequence, kv head, chunk).  Same arithmetic; the split
                                       points differ, so results agree to fp32 rounding.  Measured on B200
                                       it wins only on one-kv-head shards of large batches (by 1-2 %), so
                                       it is never chosen automatically.  Ignored where its conditions do
                                       not hold. */
#define MVL_DECODE_OVERLAP_PREV 512u /* 16-bit caches: launch the decode kernel with a programmatic
                                       dependency (PDL) on the kernel in front of it on the stream, so
                                       that its start-up — context lengths, block tables, the first K/V
                                       tiles — overlaps that kernel's tail; q, the new token's K/V and
                                       every write wait for that kernel to complete.  A PROMISE by the
                                       caller: the kernel in front writes no K/V-cache row this call
                                       reads other than each sequence's newest token (context_lens - 1).
                                       True between the layers of a decode step and between decode
                                       steps (reference qwen3.py:104-111: norm, RoPE and the store of the
                                       new token sit in front of attention); NOT true right behind a
                                       store_kvcache of older rows, e.g. a prefill's. */

/* Library / ABI identification.  No reference counterpart (the reference has no native layer). */
int mvl_abi_version(void);

/* Human-readable text for an mvl_status. */
const char* mvl_strerror(int status);

/* Text of the last CUDA error seen by the calling thread inside this library ("" if none). */
const char* mvl_last_cuda_error(void);

/*
 * mvl_store_kvcache — slot-mapped scatter of new K/V rows into the paged cache.
 * Replaces: store_kvcache() + store_kvcache_kernel, reference attention.py:67-108 and :7-64.
 *
 *   for t in [0, num_tokens):  s = slot_mapping[t];  if s < 0 (reference: == -1) skip;
 *     k_cache.view(-1, KVH, D)[s] = key[t];  v_cache.view(-1, KVH, D)[s] = value[t]
 *
 * key/value: (num_tokens, num_kv_heads, head_dim), head dims contiguous, token stride given in
 * ELEMENTS (so the q/k/v split views of reference qwen3.py:86-101 need no .contiguous() copy).
 * slot_mapping: int32 or int64 (engine uses int64, model_runner.py:252).  num_slots =
 * num_blocks * block_size; slots >= num_slots are skipped (memory safety; the reference would
 * write out of bounds).  Bit-exact copy; cache bytes of untouched slots are not written.
 */
int mvl_store_kvcache(const void* key, const void* value, void* k_cache, void* v_cache,
                      const void* slot_mapping, int slot_is_i64, int64_t num_tokens,
                      int num_kv_heads, int head_dim, int64_t key_token_stride,
                      int64_t value_token_stride, int64_t num_slots, int elem_dtype,
                      void* stream);

/*
 * Split-KV planning.  A sequence's tokens are cut into chunks of `chunk_tokens`; each
 * (sequence, kv head, chunk) is one CTA.  mvl_paged_decode_default_chunk() returns the chunk the
 * library picks when the caller passes chunk_tokens = 0; it depends only on the arguments and the
 * SM count of the current device.  Returns a negative mvl_status on bad arguments.
 * No reference counterpart (the Triton kernel walks the whole context in one program,
 * attention.py:323-326).
 */
int mvl_paged_decode_default_chunk(int64_t batch, int num_heads, int num_kv_heads, int head_dim,
                                   int64_t max_num_blocks, int block_size, int elem_dtype);

/* Bytes of scratch mvl_paged_decode needs for these sizes (0 if one chunk covers every row). */
size_t mvl_paged_decode_workspace_bytes(int64_t batch, int num_heads, int num_kv_heads,
                                        int head_dim, int64_t max_num_blocks, int block_size,
                                        int elem_dtype, int chunk_tokens);

/*
 * mvl_paged_decode — block-table GQA decode attention over the paged cache, one query per row.
 * Replaces: paged_attention_decode() + paged_attention_decode_kernel, reference
 * attention.py:418-473 and :283-415 (same as benchmark_decoding.py:6-133).
 *
 *   out[b, h, :] = softmax_j( scale * <query[b,h,:], K[b,j,h/G,:]> ) @ V[b,:,h/G,:],
 *   j in [0, context_lens[b]),  G = num_heads / num_kv_heads,
 *   K[b,j] = k_cache[block_tables[b, j / block_size], j % block_size]      (attention.py:347-361)
 *
 * query/out: (batch, num_heads, head_dim) contiguous, elem_dtype.  block_tables: int32,
 * row stride block_table_stride (elements), entries past ceil(ctx/block_size) are never read.
 * context_lens: int32 or int64.  context_lens[b] <= 0 yields a zero row (the reference yields
 * NaN, attention.py:411; such rows are CUDA-graph padding, model_runner.py:283).
 * context_lens[b] is clamped to max_num_blocks * block_size.  Scores and softmax run in fp32;
 * one rounding to elem_dtype on store.
 * workspace: >= mvl_paged_decode_workspace_bytes(...) bytes, 16-byte aligned; may be NULL when
 * that size is 0.  Calls sharing one workspace must be ordered on the stream.
 */
int mvl_paged_decode(void* out, const void* query, const void* k_cache, const void* v_cache,
                     const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                     int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                     int num_heads, int num_kv_heads, int head_dim, int block_size,
                     int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                     size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream);

/*
 * mvl_paged_decode_peers — mvl_paged_decode whose result stores also perform the all-gather of a
 * KV-head-sharded run.  Rank r of N computes num_heads = total_heads / N local heads and writes
 * them at head offset head_offset into EVERY destination out_peers[0..num_peers): each is a
 * (batch, total_heads, head_dim) tensor, normally the peer-mapped (NVLink P2P) result buffer of
 * one rank of the node, so after the kernels of all ranks have run and the ranks have met in a
 * barrier every rank holds all heads — no separate collective, no head-major -> batch-major copy.
 * The reference has no such step (attention is followed by row-parallel o_proj + all-reduce,
 * linear.py:216-221); this is the "all-gather where o_proj needs full heads" of BASELINE.json's
 * north_star, fused into the producing kernels.  num_peers <= 8.  With num_peers = 1,
 * total_heads = num_heads, head_offset = 0 it is exactly mvl_paged_decode.
 */
int mvl_paged_decode_peers(void* const* out_peers, int num_peers, int total_heads, int head_offset,
                           const void* query, const void* k_cache, const void* v_cache,
                           const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                           int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                           int num_heads, int num_kv_heads, int head_dim, int block_size,
                           int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                           size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream);

/*
 * mvl_flash_prefill — varlen causal GQA flash attention over the NEW tokens of each sequence.
 * Replaces: flash_attention_prefill() + flash_attention_varlen_kernel, reference
 * attention.py:212-280 and :111-209.
 *   out[i, h, :] = softmax_{j <= i, j in i's sequence}( scale * <q[i,h,:], k[j,h/G,:]> ) @ v[j,h/G,:]
 * q (T, H, D), k and v (T, KVH, D): head dims dense, token strides in elements; out (T, H, D)
 * contiguous.  cu_seqlens (num_seqs + 1) int32 on the device delimits the packed sequences
 * (attention.py:136-138); max_seqlen >= the longest one sizes the grid (the reference reads it
 * back from the device each call, :263-264; the engine already has it, model_runner.py:219).
 * total_tokens = T, the packed tensors' first dimension (bounds the TMA loads: rows past it read
 * as zero).  Like the reference it attends over the new tokens only, not over cached-prefix pages
 * (mvl_flash_prefill_paged does).  fp32 tensors run on a CUDA-core kernel (fp32 products and
 * accumulation: the 1e-4 the reference's fp32 default asks for), 16-bit ones on the tensor cores:
 * Kernels: two on tcgen05 (TMA tiles, both products on the 5th-generation tensor cores, accumulators
 * and probabilities in tensor memory) — a pair kernel (256 queries per CTA, two tiles in ping-pong,
 * taken from max_seqlen >= 2048) and a one-tile kernel (64 < max_seqlen < 2048) — and an mma.sync
 * kernel for batches of short sequences and non-positive scales.  flags force one of them (the
 * tcgen05 ones still need scale > 0): measurement and test switches, results agree to rounding.
 */
#define MVL_PREFILL_MMA_SYNC 1u
#define MVL_PREFILL_ONE_TILE 2u
#define MVL_PREFILL_PAIR 4u
int mvl_flash_prefill(void* out, const void* q, const void* k, const void* v, int64_t q_token_stride,
                      int64_t k_token_stride, int64_t v_token_stride, int64_t total_tokens,
                      const int32_t* cu_seqlens, int num_seqs, int max_seqlen, int num_heads,
                      int num_kv_heads, int head_dim, float scale, int elem_dtype, unsigned flags,
                      void* stream);

/*
 * mvl_flash_prefill_paged — prefill attention that also reads cached-prefix pages: what reference
 * Attention.forward is handed when the block manager found a cached prefix (model_runner.py:209-215
 * passes block_tables and cu_seqlens_k > cu_seqlens_q) but does not do (attention.py:514-525 attends
 * over the new tokens only).  Sequence s has context_lens[s] keys in the paged caches —
 * a cached prefix followed by its n_s = cu_seqlens_q[s+1] - cu_seqlens_q[s] new tokens, whose K/V the
 * caller has already stored (mvl_store_kvcache / mvl_qk_norm_rope_store) — and query i of the
 * sequence, at position context_lens[s] - n_s + i, sees keys 0 .. context_lens[s] - n_s + i:
 *   out[cu_q[s] + i, h, :] = softmax_j( scale * <q[cu_q[s] + i, h], K_s[j, h/G]> ) @ V_s[j, h/G]
 * with K_s[j] = k_cache.view(-1, KVH, D)[block_tables[s, j / block_size] * block_size + j % block_size].
 * With no prefix (context_lens[s] == n_s) it equals mvl_flash_prefill on the same tokens.  It also
 * serves chunked prefill (n_s new tokens of a longer prompt per call).
 *   q (total_q_tokens, H, D) with a token stride; out (total_q_tokens, H, D) contiguous; caches
 *   (num_blocks, block_size, KVH, D) contiguous, block_size a power of two >= 16; block_tables int32,
 *   row stride block_table_stride, entries past a sequence's context unused (may be -1).
 * 16-bit dtypes: tcgen05 kernels only, scale > 0 (MVL_ERR_SHAPE otherwise); fp32: the CUDA-core kernel,
 * any scale.  The pair kernel whenever
 * max_seqlen_q > 128; flags as mvl_flash_prefill (MVL_PREFILL_ONE_TILE / MVL_PREFILL_PAIR).  Page rows past a sequence's context inside its last
 * page are multiplied by zero probabilities: the cache must hold finite values there (allocate it
 * zeroed, as the engine does).
 */
int mvl_flash_prefill_paged(void* out, const void* q, int64_t q_token_stride, int64_t total_q_tokens,
                            const int32_t* cu_seqlens_q, const void* k_cache, const void* v_cache,
                            const int32_t* block_tables, int64_t block_table_stride,
                            const void* context_lens, int ctx_is_i64, int num_seqs, int max_seqlen_q,
                            int num_heads, int num_kv_heads, int head_dim, int block_size,
                            int64_t num_blocks, float scale, int elem_dtype, unsigned flags, void* stream);

/*
 * mvl_qk_norm_rope_store — q/k RMSNorm + rotary embedding + KV-cache write in one pass over the
 * projections' output.  Replaces, per layer, reference models/qwen3.py:104-108 (rms_norm on q and k,
 * layers/layernorm.py:13-15: x / sqrt(mean(x^2) + eps) * gamma; rotary_emb, layers/
 * rotary_embedding.py:74-83 with the half-split rotation of :4-23) followed by the store of
 * layers/attention.py:498-510.
 *   q (N, H, D), k and v (N, KVH, D): head dims dense, token strides in elements (views of the fused
 *   QKV output are fine).  q_out (N, H, D) contiguous receives the normalised, rotated q (may be q
 *   itself when q is contiguous).  k (normalised, rotated) and v (as is) go to
 *   cache.view(-1, KVH, D)[slot_mapping[t]]; slot -1 skips the token; slot_mapping NULL = no cache
 *   write.  gamma (D) fp32 or NULL (no RMSNorm: the reference skips it for qkv_bias models,
 *   qwen3.py:104).  cos_sin_cache (max_position, D) fp32 laid out [cos(D/2) | sin(D/2)] per
 *   position (rotary_embedding.py:62-68) or NULL (no rotation); positions (N,) int32/int64.
 * Arithmetic in fp32, one rounding to elem_dtype on store.
 */
int mvl_qk_norm_rope_store(void* q_out, const void* q, const void* k, const void* v,
                           int64_t q_token_stride, int64_t k_token_stride, int64_t v_token_stride,
                           const float* gamma, float eps, const float* cos_sin_cache,
                           int64_t max_position, const void* positions, int pos_is_i64,
                           void* k_cache, void* v_cache, const void* slot_mapping, int slot_is_i64,
                           int64_t num_slots, int64_t num_tokens, int num_heads, int num_kv_heads,
                           int head_dim, int elem_dtype, void* stream);

/*
 * mvl_decode_step — store + decode of one decode step in ONE kernel launch (plus the combine unless
 * MVL_DECODE_FUSED_COMBINE): what reference Attention.forward does with two Triton launches
 * (attention.py:498-510 store_kvcache, then :544-555 paged_attention_decode).  key/value are the
 * step's new rows, (batch, num_kv_heads, head_dim) with token strides in elements; slot_mapping[b]
 * must be the slot of token context_lens[b]-1 of sequence b (what reference
 * model_runner.py:238-239 produces) or -1.  The CTA that streams the chunk holding that token
 * writes the row first, so no separate store kernel runs; rows with context_lens[b] <= 0 store
 * nothing.  slot_mapping and context_lens share one index dtype.  Other arguments as
 * mvl_paged_decode.
 */
int mvl_decode_step(void* out, const void* query, const void* key, const void* value, void* k_cache,
                    void* v_cache, const void* slot_mapping, const int32_t* block_tables,
                    const void* context_lens, int idx_is_i64, int64_t batch, int64_t max_num_blocks,
                    int64_t block_table_stride, int64_t key_token_stride, int64_t value_token_stride,
                    int num_heads, int num_kv_heads, int head_dim, int block_size,
                    int64_t num_blocks, float scale, int elem_dtype, void* workspace,
                    size_t workspace_bytes, int chunk_tokens, unsigned flags, void* stream);

/* mvl_decode_step with the result stores of mvl_paged_decode_peers (head-sharded step whose stores
 * perform the all-gather).  No reference counterpart. */
int mvl_decode_step_peers(void* const* out_peers, int num_peers, int total_heads, int head_offset,
                          const void* query, const void* key, const void* value, void* k_cache,
                          void* v_cache, const void* slot_mapping, const int32_t* block_tables,
                          const void* context_lens, int idx_is_i64, int64_t batch,
                          int64_t max_num_blocks, int64_t block_table_stride,
                          int64_t key_token_stride, int64_t value_token_stride, int num_heads,
                          int num_kv_heads, int head_dim, int block_size, int64_t num_blocks,
                          float scale, int elem_dtype, void* workspace, size_t workspace_bytes,
                          int chunk_tokens, unsigned flags, void* stream);

/*
 * mvl_allgather_heads — NCCL all-gather of head-sharded attention outputs into the unsharded layout.
 * Rank r passes its (batch, heads_local, head_dim) result; every rank receives (batch,
 * world_size * heads_local, head_dim) in out_full, rank r's heads at head offset r * heads_local —
 * where the reference's tensor-parallel layout puts them (layers/linear.py:163-165; the reference
 * itself never gathers, it all-reduces behind o_proj, linear.py:216-221).  `nccl_comm` is the
 * caller's ncclComm_t passed as void*; `staging` is caller-owned scratch of world_size * batch *
 * heads_local * head_dim elements.  One ncclAllGather plus one layout kernel on `stream`.
 * The library does not link NCCL: it calls the ncclAllGather of the libnccl.so.2 already loaded in
 * the process (the one that made the communicator).  world_size == 1 copies out_local to out_full.
 */
int mvl_allgather_heads(void* out_full, void* staging, const void* out_local, int64_t batch,
                        int heads_local, int head_dim, int elem_dtype, int world_size, void* nccl_comm,
                        void* stream);

/*
 * mvl_decode_step_peers_sync / mvl_peer_wait — the peer-store gather with its own completion
 * protocol, for use between steps WITHOUT a barrier collective.  No reference counterpart (the
 * reference gathers nothing: it all-reduces after a row-parallel o_proj, layers/linear.py:216-221).
 *
 * mvl_paged_decode_peers / mvl_decode_step_peers leave two things to the caller: telling when every
 * rank's rows have landed (a barrier after the launch), and making sure no rank starts its next step
 * while another still reads this step's rows (the next step's stores go into the same buffer).  The
 * _sync form settles both:
 *   - each rank owns a CONTROL BLOCK of 64 bytes (16 x uint32, zero before the first step, peer-mapped
 *     like the result buffers); sync_peers[r] is rank r's block as mapped on this device.  When the
 *     launch's last rows are stored, this rank's step count goes up by one and is written to slot
 *     [my_rank] of every rank's block (from the persistent kernel's own epilogue where that kernel
 *     runs, else by a one-thread launch behind the decode kernels).
 *   - mvl_peer_wait(own block) enqueues a one-warp kernel that returns once all num_peers slots have
 *     reached this rank's own step count, i.e. once every rank's rows of this step are in this rank's
 *     buffer.  Put it on the stream in front of whatever consumes the buffer.  Every rank must make
 *     the same sequence of calls.  If a rank never arrives within timeout_ms (<= 0: 10 s) the wait
 *     gives up and writes 1 + (rank) into word [10] of the block instead of hanging the GPU.
 *   - use TWO result buffers per rank and alternate them step by step (pass the step's set in
 *     out_peers): a rank that has passed the wait of step s may store step s+1 while a slower rank
 *     still reads step s, but nobody can reach step s+2 — the next to write step s's buffer — before
 *     its own wait of step s+1, which every rank signals only behind its step-s consumer.
 * key / value / slot_mapping all NULL = no fused KV-cache write (mvl_paged_decode_peers); otherwise
 * as mvl_decode_step_peers.
 */
int mvl_decode_step_peers_sync(void* const* out_peers, void* const* sync_peers, int num_peers, int my_rank,
                               int total_heads, int head_offset, const void* query, const void* key,
                               const void* value, void* k_cache, void* v_cache, const void* slot_mapping,
                               const int32_t* block_tables, const void* context_lens, int idx_is_i64,
                               int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                               int64_t key_token_stride, int64_t value_token_stride, int num_heads,
                               int num_kv_heads, int head_dim, int block_size, int64_t num_blocks,
                               float scale, int elem_dtype, void* workspace, size_t workspace_bytes,
                               int chunk_tokens, unsigned flags, void* stream);
int mvl_peer_wait(void* sync_local, int num_peers, int64_t timeout_ms, void* stream);

/* Diagnostics only (scripts/diag_stream_times.py): device buffer of 5 x uint64 per CTA into which the
 * persistent stream kernel writes %globaltimer at entry / piece known / first tile landed / exit, and its %smid;
 * NULL (the default) switches it off.  Process-wide, not thread-safe.  No reference counterpart. */
void mvl_debug_set_times(void* buf);

/*
 * mvl_decode_step_host — one decode-attention step of a layer driven from HOST buffers:
 * async H2D of the step's inputs (pinned host memory recommended) into caller-provided device
 * staging buffers, mvl_store_kvcache, mvl_paged_decode, async D2H of the result.  The caller
 * synchronises the stream before reading out_host.  Mirrors what reference
 * model_runner.py:231-256 (prepare_decode's H2D copies) + attention.py:493-557
 * (Attention.forward decode branch) do per layer.  slot_mapping/context_lens are int64 when
 * idx_is_i64 != 0 else int32.  Staging buffers must hold at least the host arrays' bytes.
 * Runs mvl_decode_step with MVL_DECODE_FUSED_COMBINE | flags: the first 256 KiB of the workspace
 * must be zero before the first call.
 * Zero-copy: where a *_dev pointer EQUALS its *_host pointer (pinned, device-mapped host memory) that array
 * is not copied; the kernels read it / write it across PCIe themselves as they go.  Honoured for q, for k and
 * v together (the new token's rows, read once by the fused store) and for out (rows written as they
 * complete); slots, lengths and block tables are always copied, the kernels read them many times.  Same
 * results bit for bit; on B200 it takes 0.09 ms off a 0.86 ms step (C3), because neither the q copy in
 * front of the kernel nor the result copy behind it is exposed.  (With MVL_DECODE_SHARE_PREFIX prefer q in
 * the copy: the shared-chunk kernel fetches q rows once per work item, a PCIe round trip each time.)
 */
int mvl_decode_step_host(const void* q_host, const void* k_host, const void* v_host,
                         const void* slot_mapping_host, const void* context_lens_host,
                         const int32_t* block_tables_host, void* out_host, void* q_dev,
                         void* k_dev, void* v_dev, void* slot_mapping_dev, void* context_lens_dev,
                         int32_t* block_tables_dev, void* out_dev, void* k_cache, void* v_cache,
                         int idx_is_i64, int64_t batch, int64_t max_num_blocks, int num_heads,
                         int num_kv_heads, int head_dim, int block_size, int64_t num_blocks,
                         float scale, int elem_dtype, void* workspace, size_t workspace_bytes,
                         int chunk_tokens, unsigned flags, void* stream);

/*
 * mvl_decode_step_host_overlapped — mvl_decode_step_host with the copies hidden under the kernels.
 * The batch is cut into `pieces` (1..8) runs of consecutive sequences; each piece's host-to-device
 * copies, its decode launch and the device-to-host copy of its rows go to one of the library's own
 * streams, so piece i+1's inputs travel while piece i computes and piece i's rows return while piece
 * i+1 computes (reference model_runner.py:231-256 + attention.py:493-557 are strictly serial).
 * Ordering seen by the caller is that of mvl_decode_step_host: everything starts after the work
 * already on `stream` and `stream` waits for all of it, so one synchronisation of `stream` makes
 * out_host valid.  Same arguments as mvl_decode_step_host plus `pieces`; one cudaMemcpyAsync per
 * array and piece (6 in, 1 out).
 *   - The library creates its streams and events (no device memory) on the first call on each
 *     device; make that first call outside any stream capture.
 *   - The workspace is cut into `pieces` equal regions: size it with
 *     mvl_decode_step_host_workspace_bytes() and zero it before the first call.
 *   - Calls on one device share those internal streams: issue them from one host thread at a time.
 * No reference counterpart.
 */
size_t mvl_decode_step_host_workspace_bytes(int64_t batch, int num_heads, int num_kv_heads, int head_dim,
                                            int64_t max_num_blocks, int block_size, int elem_dtype,
                                            int chunk_tokens, int pieces);
int mvl_decode_step_host_overlapped(const void* q_host, const void* k_host, const void* v_host,
                                    const void* slot_mapping_host, const void* context_lens_host,
                                    const int32_t* block_tables_host, void* out_host, void* q_dev,
                                    void* k_dev, void* v_dev, void* slot_mapping_dev,
                                    void* context_lens_dev, int32_t* block_tables_dev, void* out_dev,
                                    void* k_cache, void* v_cache, int idx_is_i64, int64_t batch,
                                    int64_t max_num_blocks, int num_heads, int num_kv_heads,
                                    int head_dim, int block_size, int64_t num_blocks, float scale,
                                    int elem_dtype, void* workspace, size_t workspace_bytes,
                                    int chunk_tokens, unsigned flags, int pieces, void* stream);

/*
 * mvl_paged_decode_static — mvl_paged_decode for callers that cannot allocate device memory.
 * This entry point has no counterpart in the reference (its Triton kernel needs no scratch).
 *
 * Same arguments and result as mvl_paged_decode, minus the workspace: the split-KV partials go
 * into an array of MVL_STATIC_SCRATCH_BYTES that is part of this library's device image.  Use it
 * when your side of the boundary has nothing to allocate from — all free HBM already belongs to
 * the KV cache, your binding holds raw pointers but no allocator, or your loop cannot take the
 * device-wide synchronisation a cudaMalloc implies.  The call itself performs no cudaMalloc,
 * cudaFree or any other allocating call.
 *
 *   - Cost: every GPU on which the library is loaded keeps those MVL_STATIC_SCRATCH_BYTES resident
 *     until the process exits, whether or not this function is ever called.
 *   - Limit: if mvl_paged_decode_workspace_bytes(...) for the given sizes exceeds
 *     MVL_STATIC_SCRATCH_BYTES the call returns MVL_ERR_LIMIT and launches nothing;
 *     mvl_static_last_error() then says how many bytes were needed.  A larger chunk_tokens
 *     (fewer chunks per sequence) or a smaller batch shrinks the need.
 *   - First call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *     that device current).  Make that first call outside any stream capture.  The address is
 *     kept until the process exits; after a cudaDeviceReset() on that device it is stale, so do
 *     not call this function on a device you have reset.
 *   - All calls on a device share the one array and it keeps whatever the previous call left;
 *     nothing in it needs to start at zero, so nothing is cleared.  Two calls that may run at the
 *     same time on one GPU (different streams, or different host threads) race on it: keep calls
 *     ordered on one stream, or order the streams with events.
 */
#define MVL_STATIC_SCRATCH_BYTES (16u << 20) /* 16 MiB */

int mvl_paged_decode_static(void* out, const void* query, const void* k_cache, const void* v_cache,
                            const int32_t* block_tables, const void* context_lens, int ctx_is_i64,
                            int64_t batch, int64_t max_num_blocks, int64_t block_table_stride,
                            int num_heads, int num_kv_heads, int head_dim, int block_size,
                            int64_t num_blocks, float scale, int elem_dtype, int chunk_tokens,
                            void* stream);

/* MVL_STATIC_SCRATCH_BYTES as seen by the built library. */
size_t mvl_static_scratch_bytes(void);

/* Text explaining the last failure of mvl_paged_decode_static on the calling thread ("" if none):
 * the size that did not fit, the failed address look-up, or the status text (and CUDA error string)
 * of the shared decode path when that refused the arguments or the launch. */
const char* mvl_static_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* MVL_ABI_H_ */
