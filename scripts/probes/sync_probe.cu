// Microbenchmark 2: (A) latency of {K x tcgen05.mma (TS, N) ; commit ; mbarrier wait} rounds and
// (B) one-way mbarrier signalling latency between two warps (ping-pong), (C) tcgen05.st x32 +
// wait::st latency.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I miraculix_b200/csrc
#include <cstdio>
#include <cuda_runtime.h>
#include "common.cuh"
using namespace mb200;

template <int K>
__global__ void __launch_bounds__(128, 1) mma_round(int N, int iters, long long *out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 32 * 1024 / 4; i += 128) ((uint32_t *)smem)[i] = 0x01010101u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) tmem_alloc(&slot, 512);
  fence_proxy_async(); tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = make_idesc_i8(128, N, 1, 1);
    const uint64_t bdesc = make_desc_k_sw128(smem_u32(smem));
    uint32_t ph = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < K; k++) mma_i8_ts(tm, tm + 384 + 8 * (k & 3), bdesc + 2 * (k & 3), idesc, 1);
        tc_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, ph);
      ph ^= 1;
    }
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) *out = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

__global__ void __launch_bounds__(64, 1) pingpong(int iters, long long *out) {
  __shared__ uint64_t b0, b1;
  if (threadIdx.x == 0) { mbar_init(&b0, 1); mbar_init(&b1, 1); fence_mbar_init(); }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t ph = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
    if (warp == 0) {
      if (lane == 0) mbar_arrive(&b0);
      mbar_wait(&b1, ph);
    } else {
      mbar_wait(&b0, ph);
      __syncwarp();
      if (lane == 0) mbar_arrive(&b1);
    }
    ph ^= 1;
  }
  long long t1 = clock64();
  if (blockIdx.x == 0 && threadIdx.x == 0) *out = t1 - t0;
}

__global__ void __launch_bounds__(128, 1) st_lat(int iters, long long *out) {
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) tmem_alloc(&slot, 512);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot + ((uint32_t)(threadIdx.x & ~31) << 16);
  uint32_t a[32];
  for (int i = 0; i < 32; i++) a[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
    tmem_st_32x32b_x32(tm + 32 * (it & 7), a);
    tmem_st_wait();
    a[it & 31] += 1;
  }
  long long t1 = clock64();
  if (blockIdx.x == 0 && threadIdx.x == 0) *out = t1 - t0;
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(slot, 512); }
}

template <int K> void runA(long long *d) {
  long long h;
  cudaFuncSetAttribute(mma_round<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
  for (int N : {16, 80}) {
    mma_round<K><<<148, 128, 48 * 1024>>>(N, 1024, d);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); exit(1); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("A: %2d MMAs(N=%2d)+commit+wait : %.1f cycles/round\n", K, N, (double)h / 1024);
  }
}
int main() {
  long long *d, h;
  cudaMalloc(&d, 8);
  runA<1>(d); runA<2>(d); runA<4>(d); runA<8>(d); runA<16>(d);
  pingpong<<<148, 64>>>(4096, d);
  cudaDeviceSynchronize(); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("B: mbarrier ping-pong one-way latency %.1f cycles\n", (double)h / 4096 / 2);
  st_lat<<<148, 128>>>(4096, d);
  cudaDeviceSynchronize(); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("C: tcgen05.st x32 + wait::st (4 warps) %.1f cycles\n", (double)h / 4096);
  return 0;
}
