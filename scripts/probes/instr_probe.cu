// Microbenchmark 3: cost (cycles, single warp, back to back) of the bookkeeping instructions of
// the MMA-issuer loop.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I miraculix_b200/csrc
#include <cstdio>
#include <cuda_runtime.h>
#include "common.cuh"
using namespace mb200;

__global__ void __launch_bounds__(32, 1) probe(int iters, long long *out) {
  __shared__ uint64_t bar[4];
  __shared__ uint32_t slot;
  if (threadIdx.x == 0) { for (int i = 0; i < 4; i++) mbar_init(&bar[i], 1); fence_mbar_init(); }
  tmem_alloc(&slot, 64);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  long long t0, t1;
  int k = 0;
  // (a) try_wait on a barrier whose phase 1 is "already complete" (parity trick) + vote
  t0 = clock64();
  for (int i = 0; i < iters; i++) mbar_wait_warp(&bar[i & 3], 1);
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0; k++;
  // (b) plain per-lane wait
  t0 = clock64();
  for (int i = 0; i < iters; i++) mbar_wait(&bar[i & 3], 1);
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0; k++;
  // (c) tcgen05.fence::after_thread_sync
  t0 = clock64();
  for (int i = 0; i < iters; i++) tc_fence_after();
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0; k++;
  // (d) elect + syncwarp (empty body with a side effect)
  t0 = clock64();
  unsigned acc = 0;
  for (int i = 0; i < iters; i++) { if (elect_one()) acc += i; __syncwarp(); }
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0 + (acc == 12345); k++;
  // (e) clock64 itself
  t0 = clock64();
  long long s = 0;
  for (int i = 0; i < iters; i++) s += clock64();
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0 + (s == 1); k++;
  // (f) elect + tcgen05.commit (no MMAs pending) + syncwarp, then wait for it
  t0 = clock64();
  uint32_t ph = 0;
  for (int i = 0; i < iters; i++) {
    if (elect_one()) tc_commit(&bar[0]);
    __syncwarp();
    mbar_wait_warp(&bar[0], ph); ph ^= 1;
  }
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0; k++;
  // (g) elect + tcgen05.commit only, never waited (arrivals just flip phases)
  t0 = clock64();
  for (int i = 0; i < iters; i++) { if (elect_one()) tc_commit(&bar[1 + (i & 1)]); __syncwarp(); }
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0; k++;
  // (h) mbarrier arrive by lane 0 + syncwarp
  t0 = clock64();
  for (int i = 0; i < iters; i++) { if (threadIdx.x == 0) mbar_arrive(&bar[3]); __syncwarp(); }
  t1 = clock64(); if (threadIdx.x == 0) out[k] = t1 - t0; k++;
  tc_fence_before(); __syncthreads(); tc_fence_after();
  tmem_dealloc(slot, 64);
}

int main() {
  long long *d, h[8];
  cudaMalloc(&d, 64);
  const int iters = 4096;
  probe<<<1, 32>>>(iters, d);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
  cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
  const char *names[] = {"try_wait(done)+vote", "try_wait(done) per lane", "tcgen05.fence::after", "elect+syncwarp",
                         "clock64", "elect+commit+syncwarp+wait", "elect+commit+syncwarp", "lane0 arrive+syncwarp"};
  for (int i = 0; i < 8; i++) printf("%-30s %.1f cycles\n", names[i], (double)h[i] / iters);
  return 0;
}
