// prefill.h — parameters shared by the two prefill kernels (prefill.cu: mma.sync; prefill_tcgen05.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace mvl {

struct PrefillParams {
  void* out;          // (T, H, D) contiguous
  const void* q;      // (T, H, D), token stride q_stride (elements)
  const void* k;      // (T, KVH, D), token stride k_stride
  const void* v;
  const int* cu_seqlens;  // (S + 1)
  long long q_stride, k_stride, v_stride;
  long long total_tokens;
  int H, KVH;
  float qscale;       // scale * log2(e)
  // paged variant (mvl_flash_prefill_paged): k / v are the paged caches, keys 0..ctx-1 of a sequence
  // are found through its block-table row, the n new queries sit at positions ctx-n..ctx-1
  const int* block_tables;  // nullptr: dense variant
  long long bt_stride;
  const void* context_lens;
  int ctx_is_i64;
  int bs_shift;             // log2(block_size)
  int piece_rows;           // min(block_size, 128): slots fetched by one TMA box
  int num_slots;            // num_blocks * block_size (also the out-of-range coordinate that reads zeros)
};

// Tensor map over a packed (tokens, heads, D) 16-bit tensor with a token stride: box = 64 elements
// (128 bytes, the swizzle span) x 1 head x box_rows tokens, 128-byte swizzle, rows past the end
// zero-filled.  Returns 0, the CUresult of a failed encode, or -1 without the driver entry point.
int make_packed_map(CUtensorMap* map, const void* base, int dtype, int D, int heads, long long tokens,
                     long long token_stride, int box_rows);

// Tensor map over q (batch, heads, D): box = 64 elements x box_heads heads x box_seqs sequences.
int make_group_q_map(CUtensorMap* map, const void* q, int dtype, int D, int heads, long long batch, int box_heads,
                     int box_seqs);

// Record a failure for mvl_last_cuda_error() and return MVL_ERR_CUDA (abi.cu).
int fail_text(const char* where, int code);
int fail_cuda(int cuda_error, const char* where);

// tcgen05 kernel (bf16 / fp16, D in {64, 128}); returns MVL_OK or MVL_ERR_CUDA (text recorded).
// Paged variant: p.block_tables != nullptr, p.k / p.v are the caches (num_slots, KVH, D).
int launch_prefill_tcgen05(const PrefillParams& p, int dtype, int head_dim, int num_seqs, int max_seqlen,
                           bool single_tile, cudaStream_t s);

// fp32 tensors: CUDA-core kernel (prefill_f32.cu), dense or paged; returns MVL_OK or MVL_ERR_CUDA.
int launch_prefill_f32(const PrefillParams& p, int head_dim, int num_seqs, int max_seqlen, cudaStream_t s);

}  // namespace mvl
