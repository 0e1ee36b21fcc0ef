// noalloc.cu -- mb200_dgemm_compressed_noalloc: the fp64 dgemm_compressed on device pointers for
// callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//
// The ordinary path keeps its centring scalars w[c] = -2 sum_k x_k B[k,c] (reference
// dgemm_compressed_cuda.cu:421-459) in scratch that belongs to the handle.  This variant keeps
// them in a file-scope __device__ array of MB200_NOALLOC_MAX_COLS doubles (8 KiB, held by every
// GPU that loads the library until the process exits).  In host code the array's name is only a
// stand-in: its device address is looked up once per device with cudaGetSymbolAddress, with that
// device current, and that pointer is what the kernels receive.
//
// The entry point calls no cudaMalloc / cudaFree / allocating routine: a memset of the scratch and
// three kinds of kernel launches on the handle's stream.  All calls share the array, which keeps
// the previous call's scalars: it is cleared at the start of every call (the scalars are
// accumulated with atomics), and two calls that overlap on one GPU (different streams / handles)
// race on it.
#include <mutex>
#include "common.cuh"
#include "handle.h"

#define MB200_NOALLOC_MAX_COLS 1024  // include/miraculix_b200.h states the same limit

// The scratch: file scope, one copy per GPU that loads the library.
static __device__ __align__(256) double g_noalloc_w[MB200_NOALLOC_MAX_COLS];

namespace mb200 {

// dgemm.cu: the fp64 streaming kernel launches (C += Z_op B into the centre-initialised C)
int dgemm_fp64_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc);

namespace {

constexpr int kMaxDev = 64;
std::mutex g_mu;
double *g_w_addr[kMaxDev] = {nullptr};

// device address of g_noalloc_w on `device`, which must be the current device
int scratch_address(int device, double **out) {
  if (device < 0 || device >= kMaxDev) return fail(-4, __func__, __LINE__, "device %d out of range", device);
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_w_addr[device]) {
    double *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_w);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(-4, __func__, __LINE__, "cudaGetSymbolAddress(scratch) on device %d: %s", device,
                  cudaGetErrorString(e));
    }
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess || at.type != cudaMemoryTypeDevice || at.device != device) {
      cudaGetLastError();
      return fail(-4, __func__, __LINE__, "scratch symbol does not resolve to memory of device %d", device);
    }
    g_w_addr[device] = p;
  }
  *out = g_w_addr[device];
  return 0;
}

// w[col] += -2 sum_k x_k B[k,col] over this block's share of k; w is the static scratch
__global__ void __launch_bounds__(256)
na_center_w_kernel(const double *__restrict__ B, long ldb, long k, const double *__restrict__ x,
                   double *__restrict__ w) {
  const int col = blockIdx.x;
  const double *b = B + (size_t)col * ldb;
  double acc = 0.0;
  for (long j = (long)blockIdx.y * 256 + threadIdx.x; j < k; j += (long)gridDim.y * 256)
    acc += (x ? x[j] : 1.0) * b[j];
  __shared__ double red[8];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double t = threadIdx.x < 8 ? red[threadIdx.x] : 0.0;
    for (int o = 4; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) atomicAdd(&w[col], -2.0 * t);
  }
}

// C[r,c] = w[c] y_r (centred) or 0
__global__ void __launch_bounds__(256)
na_init_c_kernel(double *__restrict__ C, long ldc, long m, int n, const double *__restrict__ w,
                 const double *__restrict__ y, int centered) {
  const long r = (long)blockIdx.x * 256 + threadIdx.x;
  if (r >= m) return;
  const double yr = y ? y[r] : 1.0;
  for (int c = 0; c < n; c++) C[r + (size_t)c * ldc] = centered ? w[c] * yr : 0.0;
}

}  // namespace

int dgemm_noalloc_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc) {
  double *w = nullptr;
  MB_CHECK(scratch_address(h->device, &w));
  const DgemmView v = full_view(h, trans);
  const int centered = options().centered;
  if (centered) {
    MB_CUDA(cudaMemsetAsync(w, 0, sizeof(double) * n, h->stream));
    long ysplit = ceil_div(v.K, 256L * 16);
    if (ysplit > 64) ysplit = 64;
    if (ysplit < 1) ysplit = 1;
    na_center_w_kernel<<<dim3((unsigned)n, (unsigned)ysplit), 256, 0, h->stream>>>(dB, ldb, v.K, v.x, w);
    count_launch();
    MB_LAUNCH_CHECK();
  }
  na_init_c_kernel<<<(unsigned)ceil_div(v.rows, 256), 256, 0, h->stream>>>(dC, ldc, v.rows, n, w, v.y, centered);
  count_launch();
  MB_LAUNCH_CHECK();
  return dgemm_fp64_device(h, trans, n, dB, ldb, dC, ldc);
}

}  // namespace mb200

extern "C" {

int mb200_noalloc_max_cols(void) { return MB200_NOALLOC_MAX_COLS; }

int mb200_dgemm_compressed_noalloc(int trans, void *compressed, int n, const double *dB, long ldb,
                                   double *dC, long ldc) {
  using namespace mb200;
  clear_status();
  if (n > MB200_NOALLOC_MAX_COLS)
    return fail(-2, __func__, __LINE__, "n=%d exceeds the static scratch of %d columns (MB200_NOALLOC_MAX_COLS)",
                n, MB200_NOALLOC_MAX_COLS);
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  const long m = trans ? h->snps : h->indiv, k = trans ? h->indiv : h->snps;
  if (n < 0 || ldb < k || ldc < m || (n > 0 && (!dB || !dC)))
    return fail(-2, __func__, __LINE__, "bad arguments n=%d ldb=%ld (k=%ld) ldc=%ld (m=%ld) dB=%p dC=%p", n, ldb,
                k, ldc, m, (const void *)dB, (void *)dC);
  if (n == 0) return 0;
  return dgemm_noalloc_device(h, trans ? 1 : 0, n, dB, ldb, dC, ldc);
}

}  // extern "C"
