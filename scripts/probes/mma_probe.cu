// Microbenchmark: cycles per tcgen05.mma kind::i8 (M=128, K=32) as a function of N, of where A
// comes from (shared memory descriptor vs TMEM), of how many accumulators rotate, and of how the
// issuing thread is selected (lane==0 branch vs whole warp + elect.sync).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I miraculix_b200/csrc -o mma_probe mma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "common.cuh"
using namespace mb200;

template <int TS, int NACC, int ELECT, int MMAS = 16>
__global__ void __launch_bounds__(128, 1) probe(int N, int iters, long long *out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) ((uint32_t *)smem)[i] = 0x01010101u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) tmem_alloc(&slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32 && (ELECT || threadIdx.x == 0)) {
    const uint32_t idesc = make_idesc_i8(128, N, 1, 1);
    const uint64_t adesc = make_desc_k_sw128(smem_u32(smem));
    const uint64_t bdesc = make_desc_k_sw128(smem_u32(smem + 16384));
    uint32_t ph = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      if (!ELECT || elect_one()) {
#pragma unroll
        for (int k = 0; k < MMAS; k++) {
          const uint32_t dd = tm + (k % NACC) * N;
          if (TS) mma_i8_ts(dd, tm + 384 + 8 * (k & 3), bdesc + 2 * (k & 3), idesc, 1);
          else mma_i8_ss(dd, adesc + 2 * (k & 3), bdesc + 2 * (k & 3), idesc, 1);
        }
        tc_commit(&bar);
      }
      if (ELECT) __syncwarp();
      mbar_wait(&bar, ph);
      ph ^= 1;
    }
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) *out = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

template <int TS, int NACC, int ELECT, int MMAS = 16>
void run(long long *d) {
  long long h;
  auto kern = probe<TS, NACC, ELECT, MMAS>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  const int Ns[] = {16, 32, 64, 80, 96, 128, 256};
  for (int N : Ns) {
    if (N * NACC > 384) continue;
    const int iters = 256;
    kern<<<148, 128, 80 * 1024>>>(N, iters, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%s nacc=%d %s mmas/commit=%d N=%3d  cycles/mma=%6.1f (floor %d)\n", TS ? "TS" : "SS", NACC,
           ELECT ? "elect " : "lane0 ", MMAS, N, (double)h / (iters * MMAS), N / 2);
  }
}

int main() {
  long long *d;
  cudaMalloc(&d, 8);
  run<0, 1, 1, 64>(d); run<0, 2, 1, 64>(d);  // SS, long dependent chains into 1 or 2 accumulators
  run<0, 1, 0>(d); run<0, 1, 1>(d); run<0, 4, 1>(d);
  run<1, 1, 0>(d); run<1, 1, 1>(d); run<1, 2, 1>(d); run<1, 4, 1>(d);
  return 0;
}
