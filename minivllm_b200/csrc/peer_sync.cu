// peer_sync.cu — completion signal and wait for the peer-store gather of head-sharded decode.
//
// With head sharding every rank's decode kernels store their heads straight into every rank's
// (batch, heads, head_dim) result buffer over NVLink (mvl_paged_decode_peers).  A consumer may read
// its buffer once ALL ranks have stored.  Instead of a barrier collective after the kernel, each rank
// announces "my step s is stored" by writing s into its slot of every rank's control block, and a
// consumer waits until all slots of ITS block have reached its own step count:
//   * the persistent stream kernel signals from its own epilogue (last CTA to arrive);
//   * after any other decode kernel, peer_signal_kernel (one thread) does it in stream order;
//   * peer_wait_kernel (one warp, lane r polls rank r's slot) goes in front of the consumer.
// Step counts only grow, so nothing is ever reset and a slow rank cannot miss a signal.  The result
// buffers are used in turns (two per rank): a rank that has passed the wait of step s may already
// store step s+1 into the other buffer while a slower rank still reads step s; it can only reach
// step s+2 — the first to reuse step s's buffer — after its own wait of step s+1, which every rank
// signals only after its step-s consumer in stream order.
#include "host.h"

namespace mvl {

// One warp: lane 0 bumps this rank's step count, one system-scope fence orders everything the decode
// kernels in front stored (kernel boundary) and the count before the announcements, then lane r writes
// the count into rank r's block — all ranks at once, relaxed stores behind the one fence (a release
// store per peer would pay the NVLink round trip eight times in a row).
__global__ void peer_signal_kernel(unsigned* local, unsigned* p0, unsigned* p1, unsigned* p2, unsigned* p3,
                                   unsigned* p4, unsigned* p5, unsigned* p6, unsigned* p7, int n, int my_rank) {
  unsigned* peers[8] = {p0, p1, p2, p3, p4, p5, p6, p7};
  const int lane = threadIdx.x;
  unsigned step = 0;
  if (lane == 0) {
    step = local[8] + 1u;
    local[8] = step;
  }
  step = __shfl_sync(0xffffffffu, step, 0);
  __threadfence_system();
  if (lane < n)
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(peers[lane] + my_rank), "r"(step) : "memory");
}

__global__ void peer_wait_kernel(unsigned* local, int n, unsigned long long timeout_ns) {
  // a decode kernel launched behind this wait with MVL_DECODE_OVERLAP_PREV may begin its start-up now;
  // it orders itself behind this kernel's completion with griddepcontrol.wait
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // launched with a programmatic dependency itself: the decode kernel in front must have completed
  // (its last CTA wrote this rank's step count) before the count is read
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int r = threadIdx.x;
  if (r < n) {
    const unsigned target = local[8];
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(local + r) : "memory");
      if ((int)(v - target) >= 0) break;
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > timeout_ns) {  // a rank never arrived: leave a mark instead of hanging the GPU
        local[10] = 1u + (unsigned)r;
        break;
      }
      __nanosleep(64);
    }
  }
  __syncwarp();
  __threadfence_system();
}

int launch_peer_signal(unsigned* local, unsigned* const* peers, int n, int my_rank, cudaStream_t s) {
  unsigned* q[8] = {};
  for (int r = 0; r < n && r < 8; ++r) q[r] = peers[r];
  peer_signal_kernel<<<1, 32, 0, s>>>(local, q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7], n, my_rank);
  return (int)cudaGetLastError();
}

int launch_peer_wait(unsigned* local, int n, unsigned long long timeout_ns, cudaStream_t s) {
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(32);
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, peer_wait_kernel, local, n, timeout_ns);
}

}  // namespace mvl
