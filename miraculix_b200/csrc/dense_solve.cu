// dense_solve.cu -- dense SPD solve behind the reference's potrs_solve_gpu.
//
// Replaces dense_solve (reference src/cuda/solve_cuda.cu:69-278) + its wrapper :947-951, the step
// that follows the GRM in examples/gblup/calculate_gblup.jl:138-181.  As SURVEY.md allows for
// this row, the factorisation itself is a cuSOLVER library call (Xpotrf / Xpotrs, lower
// triangle), not a hand-written kernel; what is ours is the plumbing that keeps the GRM -> solve
// chain on the device (mb200_potrs_solve_device), the deterministic log-determinant kernel, and
// the reference's status convention (0 ok, 1 failure, message printed).
#include <cusolverDn.h>
#include <math.h>
#include "../../include/miraculix_b200.h"
#include "common.cuh"
#include "handle.h"

namespace mb200 {

// logdet(A) = 2 sum_i log L_ii (reference trace_kernel, solve_cuda.cu:884-909, which sums with
// shared/global fp64 atomics; here one block, fixed order).
__global__ void __launch_bounds__(1024)
chol_logdet_kernel(const double *__restrict__ L, long n, double *__restrict__ out) {
  __shared__ double red[32];
  double acc = 0.0;
  for (long i = threadIdx.x; i < n; i += 1024) acc += 2.0 * log(L[i * (n + 1)]);
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = red[threadIdx.x];
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) *out = acc;
  }
}

#define CS_CHECK(expr)                                                                      \
  do {                                                                                      \
    cusolverStatus_t _s = (expr);                                                           \
    if (_s != CUSOLVER_STATUS_SUCCESS) {                                                    \
      rc = fail(-30 - (int)_s, __func__, __LINE__, "%s -> cusolver status %d", #expr, (int)_s); \
      goto done;                                                                            \
    }                                                                                       \
  } while (0)
#define CU_CHECK(expr)                                                                      \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      rc = fail((int)_e, __func__, __LINE__, "%s: %s", #expr, cudaGetErrorString(_e));      \
      goto done;                                                                            \
    }                                                                                       \
  } while (0)

// In place on the device: d_A (n x n, column-major = row-major for a symmetric matrix; only the
// lower triangle is read) is overwritten by its Cholesky factor, d_B (n x nrhs, ld n) by the
// solution.  d_logdet may be NULL.
int potrs_device(double *d_A, long n, double *d_B, long nrhs, double *d_logdet, cudaStream_t stream) {
  int rc = 0;
  cusolverDnHandle_t handle = nullptr;
  cusolverDnParams_t params = nullptr;
  void *work_d = nullptr, *work_h = nullptr;
  int *d_info = nullptr, h_info = 0;
  size_t wd = 0, wh = 0;
  CS_CHECK(cusolverDnCreate(&handle));
  CS_CHECK(cusolverDnSetStream(handle, stream));
  CS_CHECK(cusolverDnCreateParams(&params));
  CU_CHECK(cudaMalloc(&d_info, sizeof(int)));
  CU_CHECK(cudaMemsetAsync(d_info, 0, sizeof(int), stream));
  CS_CHECK(cusolverDnXpotrf_bufferSize(handle, params, CUBLAS_FILL_MODE_LOWER, n, CUDA_R_64F, d_A, n,
                                       CUDA_R_64F, &wd, &wh));
  if (wd) CU_CHECK(cudaMalloc(&work_d, wd));
  if (wh) { work_h = malloc(wh); if (!work_h) { rc = fail(-6, __func__, __LINE__, "out of host memory"); goto done; } }
  CS_CHECK(cusolverDnXpotrf(handle, params, CUBLAS_FILL_MODE_LOWER, n, CUDA_R_64F, d_A, n, CUDA_R_64F,
                            work_d, wd, work_h, wh, d_info));
  CU_CHECK(cudaMemcpyAsync(&h_info, d_info, sizeof(int), cudaMemcpyDeviceToHost, stream));
  CU_CHECK(cudaStreamSynchronize(stream));
  if (h_info != 0) {  // reference: prints and returns 1 (and resets the device, which we do not)
    if (h_info > 0) rc = fail(-40, __func__, __LINE__, "Cholesky factorization failed at minor %d", h_info);
    else rc = fail(-41, __func__, __LINE__, "Wrong parameter in cholesky factorization at %d entry", -h_info);
    goto done;
  }
  CS_CHECK(cusolverDnXpotrs(handle, params, CUBLAS_FILL_MODE_LOWER, n, nrhs, CUDA_R_64F, d_A, n,
                            CUDA_R_64F, d_B, n, d_info));
  if (d_logdet) {
    chol_logdet_kernel<<<1, 1024, 0, stream>>>(d_A, n, d_logdet);
    count_launch();
    CU_CHECK(cudaGetLastError());
  }
  CU_CHECK(cudaStreamSynchronize(stream));
done:
  if (work_d) cudaFree(work_d);
  if (work_h) free(work_h);
  if (d_info) cudaFree(d_info);
  if (params) cusolverDnDestroyParams(params);
  if (handle) cusolverDnDestroy(handle);
  return rc;
}

}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_potrs_solve_device(double *d_A, long n, double *d_B, long nrhs, double *d_logdet) {
  clear_status();
  if (!d_A || !d_B || n <= 0 || nrhs <= 0) return fail(-2, __func__, __LINE__, "bad arguments");
  return potrs_device(d_A, n, d_B, nrhs, d_logdet, 0);
}

void potrs_solve_gpu(double *A, unsigned int input_size, double *B, unsigned int rhs_cols, double *X,
                     double *logdet, int oversubscribe, int *status) {
  clear_status();
  int rc = 0;
  double *d_A = nullptr, *d_B = nullptr, *d_ld = nullptr;
  const size_t n = input_size, nrhs = rhs_cols;
  if (status) *status = 1;
  if (!A || !B || !X || n == 0 || nrhs == 0 || (oversubscribe != 0 && oversubscribe != 1)) {
    fail(-2, __func__, __LINE__, "bad arguments");
    return;
  }
  if (select_device() < 0) return;
  // oversubscribe = 1: managed memory so that A may exceed the device (solve_cuda.cu:114-122)
  CU_CHECK(oversubscribe ? cudaMallocManaged(&d_A, sizeof(double) * n * n) : cudaMalloc(&d_A, sizeof(double) * n * n));
  CU_CHECK(cudaMalloc(&d_B, sizeof(double) * n * nrhs));
  CU_CHECK(cudaMalloc(&d_ld, sizeof(double)));
  CU_CHECK(cudaMemcpy(d_A, A, sizeof(double) * n * n, cudaMemcpyHostToDevice));
  CU_CHECK(cudaMemcpy(d_B, B, sizeof(double) * n * nrhs, cudaMemcpyHostToDevice));
  rc = potrs_device(d_A, (long)n, d_B, (long)nrhs, logdet ? d_ld : nullptr, 0);
  if (rc) goto done;
  if (logdet) CU_CHECK(cudaMemcpy(logdet, d_ld, sizeof(double), cudaMemcpyDeviceToHost));
  CU_CHECK(cudaMemcpy(X, d_B, sizeof(double) * n * nrhs, cudaMemcpyDeviceToHost));
done:
  cudaFree(d_A);
  cudaFree(d_B);
  cudaFree(d_ld);
  if (status) *status = rc ? 1 : 0;
}

}  // extern "C"
