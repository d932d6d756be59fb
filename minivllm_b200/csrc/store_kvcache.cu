// store_kvcache.cu — slot-mapped scatter of new K/V rows into the paged cache (sm_100a).
//
// Replaces the reference's Triton store_kvcache_kernel (reference
// src/myvllm/layers/attention.py:7-64): for every token t with slot s = slot_mapping[t] != -1,
// cache.view(-1, KVH, D)[s] = x[t].  The flat slot address does not depend on block_size
// (block_idx*block_size + block_offset == s, attention.py:33-34,52-56).
//
// Mapping: the reference launches one 4-warp program per (token, kv head) moving 2 x 256 B.  Here a
// thread moves one 16-byte chunk of K and the matching chunk of V (128-bit LDG/STG), consecutive
// threads cover consecutive chunks of a token row, and the grid is sized from the row count with a
// grid-stride loop, so a decode step (a few hundred rows) is one small wave and a prefill (many
// thousand rows) streams at copy bandwidth.  Pure byte movement: bit-exact for any element type.
#include "common.cuh"
#include "host.h"

namespace mvl {

template <typename SlotT>
__global__ void __launch_bounds__(256)
store_kvcache_kernel(const uint4* __restrict__ key, const uint4* __restrict__ value,
                     uint4* __restrict__ k_cache, uint4* __restrict__ v_cache,
                     const SlotT* __restrict__ slot_mapping, long long num_tokens,
                     int chunks_per_row,          // KVH * D * es / 16
                     long long key_stride16,      // token stride of key in 16-byte units
                     long long value_stride16, long long num_slots) {
  const long long total = num_tokens * chunks_per_row;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / chunks_per_row;
    const int ch = (int)(i - t * chunks_per_row);
    const long long s = (long long)slot_mapping[t];
    if (s < 0 || s >= num_slots) continue;  // -1: token has no slot (attention.py:29-30)
    const uint4 kv = __ldg(key + t * key_stride16 + ch);
    const uint4 vv = __ldg(value + t * value_stride16 + ch);
    k_cache[s * chunks_per_row + ch] = kv;
    v_cache[s * chunks_per_row + ch] = vv;
  }
}

int launch_store_kvcache(const void* key, const void* value, void* k_cache, void* v_cache,
                         const void* slot_mapping, int slot_is_i64, long long num_tokens,
                         int chunks_per_row, long long key_stride16, long long value_stride16,
                         long long num_slots, int num_sms, cudaStream_t stream) {
  const long long total = num_tokens * chunks_per_row;
  if (total == 0) return 0;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms * 8;  // 8 resident 256-thread CTAs per SM
  if (blocks > cap) blocks = cap;
  if (slot_is_i64)
    store_kvcache_kernel<long long><<<(unsigned)blocks, 256, 0, stream>>>(
        (const uint4*)key, (const uint4*)value, (uint4*)k_cache, (uint4*)v_cache,
        (const long long*)slot_mapping, num_tokens, chunks_per_row, key_stride16, value_stride16,
        num_slots);
  else
    store_kvcache_kernel<int><<<(unsigned)blocks, 256, 0, stream>>>(
        (const uint4*)key, (const uint4*)value, (uint4*)k_cache, (uint4*)v_cache,
        (const int*)slot_mapping, num_tokens, chunks_per_row, key_stride16, value_stride16,
        num_slots);
  return 0;
}

}  // namespace mvl
