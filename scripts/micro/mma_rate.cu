// Microbenchmark: issue rate of back-to-back tcgen05.mma (M=128, K=16, bf16) from one thread.
//   variant 0: SS (A and B from shared memory), N = 128      variant 1: TS (A from TMEM), N = 128
//   variant 4/5: SS alternating 2/4 accumulators; 6: no accumulate
//   variant 2: SS, N = 256                                    variant 3: SS N=128 with 8 warps doing tcgen05.ld
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../minivllm_b200/csrc mma_rate.cu -o mma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "common.cuh"
#include "tcgen05.cuh"
using namespace mvl;

template <int variant>
__global__ void __launch_bounds__(384, 1) k(int iters, long long* out) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = base, sB = base + 32768, bar = base + 131072, slot = bar + 8;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 8) { tc5::tmem_alloc(slot, 512); tc5::tmem_relinquish(); }
  if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
  tc5::fence_before_sync(); __syncthreads(); tc5::fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  const int n = variant == 2 ? 256 : 128;
  const uint32_t idesc = tc5::instr_desc_f16(1, 0, 128, n);
  if (tid == 256) {
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
        if (variant == 1)
          tc5::mma_ts(tmem + 256, tmem + ks * 8, tc5::smem_desc_sw128(sB + off, 16, 1024), idesc, 1);
        else if (variant == 4)   // alternate between two independent accumulators
          tc5::mma_ss(tmem + 128 * (ks & 1), tc5::smem_desc_sw128(sA + off, 16, 1024),
                      tc5::smem_desc_sw128(sB + off, 16, 1024), idesc, 1);
        else if (variant == 5)   // four independent accumulators
          tc5::mma_ss(tmem + 128 * (ks & 3), tc5::smem_desc_sw128(sA + off, 16, 1024),
                      tc5::smem_desc_sw128(sB + off, 16, 1024), idesc, 1);
        else if (variant == 6)   // no accumulation (overwrite)
          tc5::mma_ss(tmem, tc5::smem_desc_sw128(sA + off, 16, 1024),
                      tc5::smem_desc_sw128(sB + off, 16, 1024), idesc, 0);
        else
          tc5::mma_ss(tmem + 256 * (it & 1) * (variant != 2), tc5::smem_desc_sw128(sA + off, 16, 1024),
                      tc5::smem_desc_sw128(sB + off, 16, 1024), idesc, 1);
      }
    }
    long long t1 = clock64();
    tc5::commit(bar);
    tc5::mbar_wait_bounded(bar, 0);
    long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t0;
  } else if (warp < 8 && variant == 3) {
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + 128 * (warp >> 2);
    uint32_t acc = 0;
    for (int it = 0; it < iters * 2; ++it) {
      uint32_t r[32];
      tc5::tmem_ld32(lane_base + (it & 3) * 32, r);
      tc5::tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 32; ++i) acc += r[i];
    }
    if (acc == 0x12345678) out[2] = acc;
  }
  tc5::fence_before_sync(); __syncthreads();
  if (warp == 8) tc5::tmem_dealloc(tmem, 512);
}

template <int V>
void run(long long* out) {
  cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 140000);
  for (int rep = 0; rep < 2; ++rep) { k<V><<<1, 384, 140000>>>(64, out); cudaDeviceSynchronize(); }
  printf("variant %d: issue %lld clk, done %lld clk for 512 MMAs -> %.1f / %.1f clk per MMA (%s)\n", V, out[0], out[1],
         out[0] / 512.0, out[1] / 512.0, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* out; cudaMallocManaged(&out, 64);
  run<0>(out), run<1>(out), run<2>(out), run<3>(out), run<4>(out), run<5>(out), run<6>(out);
  // all SMs busy (power / clock effects)
  k<0><<<148, 384, 140000>>>(2000, out); cudaDeviceSynchronize();
  printf("148 CTAs variant 0: %.1f clk per MMA\n", out[1] / 16000.0);
  return 0;
}
