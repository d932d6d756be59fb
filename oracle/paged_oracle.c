/*
 * paged_oracle.c — plain-C CPU restatement of the paged-KV hot path.  TEST INFRASTRUCTURE ONLY:
 * loaded (ctypes) by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by the
 * product library.  Built by oracle/Makefile into oracle/_build/libpaged_oracle.so.
 *
 * What it restates
 *   oracle_store_kvcache   reference src/myvllm/layers/attention.py:27-64 (Triton store kernel):
 *                          slot == -1 skips the token (:29-30); row address = slot * KVH * D (:52-56).
 *   oracle_paged_decode    reference src/myvllm/layers/attention.py:303-415 (Triton decode kernel)
 *                          and benchmark_decoding.py:136-189 (PyTorch version): kv head =
 *                          h / (H / KVH) (:307), token j lives at cache[table[b][j / BS]][j % BS]
 *                          (:347-361), softmax over exactly context_len tokens (:370), fp32/bf16/fp16
 *                          inputs.  Evaluated exactly: scores, a global max and the weighted sum in
 *                          double precision (no online rescaling), one float result per element.
 * Parity pinning: checked against the golden vectors produced by the reference's own
 * decode_torch_optimized (tests/golden/, tests/test_oracle_golden.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { ORACLE_F32 = 0, ORACLE_F16 = 1, ORACLE_BF16 = 2 };

static float half_to_float(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 0x1fu, man = h & 0x3ffu, bits;
  if (exp == 0) {
    if (man == 0) {
      bits = sign;
    } else { /* subnormal: normalise */
      int e = -1;
      do { man <<= 1; ++e; } while (!(man & 0x400u));
      bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
    }
  } else if (exp == 31) {
    bits = sign | 0x7f800000u | (man << 13);
  } else {
    bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
  }
  float f;
  memcpy(&f, &bits, 4);
  return f;
}

static inline double load_elem(const void* base, int64_t idx, int dtype) {
  if (dtype == ORACLE_F32) return (double)((const float*)base)[idx];
  uint16_t raw = ((const uint16_t*)base)[idx];
  if (dtype == ORACLE_BF16) {
    uint32_t bits = (uint32_t)raw << 16;
    float f;
    memcpy(&f, &bits, 4);
    return (double)f;
  }
  return (double)half_to_float(raw);
}

/* Byte-exact scatter of rows; slots are int64, -1 = skip.  row_bytes = KVH * D * elem size. */
void oracle_store_kvcache(const void* key, const void* value, void* k_cache, void* v_cache,
                          const int64_t* slot_mapping, int64_t num_tokens, int64_t row_bytes) {
  for (int64_t t = 0; t < num_tokens; ++t) {
    int64_t s = slot_mapping[t];
    if (s == -1) continue;
    memcpy((char*)k_cache + s * row_bytes, (const char*)key + t * row_bytes, (size_t)row_bytes);
    memcpy((char*)v_cache + s * row_bytes, (const char*)value + t * row_bytes, (size_t)row_bytes);
  }
}

/* out: float32 (B, H, D).  context_lens: int64.  Rows with context_len <= 0 are zeros. */
int oracle_paged_decode(float* out, const void* q, const void* k_cache, const void* v_cache,
                        const int32_t* block_tables, const int64_t* context_lens, int64_t batch,
                        int64_t max_num_blocks, int num_heads, int num_kv_heads, int head_dim,
                        int block_size, double scale, int dtype) {
  const int group = num_heads / num_kv_heads;
  int failed = 0;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
  for (int64_t b = 0; b < batch; ++b) {
    for (int h = 0; h < num_heads; ++h) {
      float* o = out + (b * num_heads + h) * head_dim;
      const int64_t n = context_lens[b];
      for (int d = 0; d < head_dim; ++d) o[d] = 0.0f;
      if (n <= 0) continue;
      const int kvh = h / group;
      double* s = (double*)malloc(sizeof(double) * (size_t)n);
      double* acc = (double*)calloc((size_t)head_dim, sizeof(double));
      double* qd = (double*)malloc(sizeof(double) * (size_t)head_dim);
      if (!s || !acc || !qd) {
        failed = 1;
        free(s); free(acc); free(qd);
        continue;
      }
      for (int d = 0; d < head_dim; ++d) qd[d] = load_elem(q, (b * num_heads + h) * head_dim + d, dtype);
      double mx = -INFINITY;
      for (int64_t j = 0; j < n; ++j) {
        const int64_t blk = block_tables[b * max_num_blocks + j / block_size];
        const int64_t row = ((blk * block_size + j % block_size) * num_kv_heads + kvh) * head_dim;
        double dot = 0.0;
        for (int d = 0; d < head_dim; ++d) dot += qd[d] * load_elem(k_cache, row + d, dtype);
        s[j] = dot * scale;
        if (s[j] > mx) mx = s[j];
      }
      double denom = 0.0;
      for (int64_t j = 0; j < n; ++j) {
        const double p = exp(s[j] - mx);
        denom += p;
        const int64_t blk = block_tables[b * max_num_blocks + j / block_size];
        const int64_t row = ((blk * block_size + j % block_size) * num_kv_heads + kvh) * head_dim;
        for (int d = 0; d < head_dim; ++d) acc[d] += p * load_elem(v_cache, row + d, dtype);
      }
      for (int d = 0; d < head_dim; ++d) o[d] = (float)(acc[d] / denom);
      free(s); free(acc); free(qd);
    }
  }
  return failed;
}
