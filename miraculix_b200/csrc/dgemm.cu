// dgemm.cu -- dgemm_compressed on the device.
//
// Replaces dgemm_compressed_gpu (reference src/cuda/dgemm_compressed_cuda.cu:218-489): CUTLASS
// SIMT u2 x f64 GEMM + cublasDgeam transpose + cublasDgemv + n x cublasDaxpy + per-call
// cublasCreate / memset / malloc.  Here:
//   center_w_kernel   w[c] = -2 sum_k x_k B[k,c]       (x = f for 'N', 1 for 'T')   (:426-437)
//   init_c_kernel     C[r,c] = w[c] * y_r               (y = 1 for 'N', f for 'T')   (:449-456)
//   main kernel       C += Z_op * B                     (:328-385, micro-kernel
//                                                        dgemm_compressed_cuda.h:253-268)
// Two main kernels exist: the fp64 SIMT streaming kernel below (mode 0) and the exact int8-slice
// tcgen05 kernel in dgemm_tc.cu (mode 1).  No cuBLAS, no per-call allocation.
#include "common.cuh"
#include "handle.h"

namespace mb200 {

int dgemm_tc_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc);
bool dgemm_tc_available();

// ---------------------------------------------------------------------------------------------
// centring
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
center_w_kernel(const double *__restrict__ B, long ldb, long k, const double *__restrict__ x,
                double *__restrict__ w, const int *__restrict__ guard) {
  if (guard && *guard == 0) return;  // precision fallback not needed (dgemm_tc.cu)
  const int col = blockIdx.x;
  const double *b = B + (size_t)col * ldb;
  double acc = 0.0;
  for (long j = (long)blockIdx.y * 256 + threadIdx.x; j < k; j += (long)gridDim.y * 256)
    acc += (x ? x[j] : 1.0) * b[j];
  __shared__ double red[8];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; i++) t += red[i];
    atomicAdd(&w[col], -2.0 * t);
  }
}
__global__ void __launch_bounds__(256)
init_c_kernel(double *__restrict__ C, long ldc, long m, int n, const double *__restrict__ w,
              const double *__restrict__ y, int centered, int accumulate,
              const int *__restrict__ guard) {
  if (guard && *guard == 0) return;
  const long r = (long)blockIdx.x * 256 + threadIdx.x;
  if (r >= m) return;
  const double yr = y ? y[r] : 1.0;
  for (int c = 0; c < n; c++) {
    const double v = centered ? w[c] * yr : 0.0;
    double *dst = &C[r + (size_t)c * ldc];
    *dst = accumulate ? *dst + v : v;
  }
}

int dgemm_init_c(Handle *h, const DgemmView &v, int n, double *dC, long ldc, int centered,
                 const double *w, const int *guard);

DgemmView full_view(const Handle *h, int trans) {
  DgemmView v;
  v.trans = trans;
  v.Z = trans ? h->d_snp_major : h->d_indiv_major;
  v.pitch = trans ? h->pitch_s : h->pitch_i;
  v.rows = trans ? h->snps : h->indiv;
  v.K = trans ? h->indiv : h->snps;
  v.x = trans ? nullptr : h->d_f;
  v.y = trans ? h->d_f : nullptr;
  v.rowsum_max = h->rowsum_max[trans ? 1 : 0];
  return v;
}
// SNP range [begin, begin + count): for 'N' a K range (begin must be a multiple of 512 SNPs =
// 128 packed bytes), for 'T' a row range.  The full-row bound on the row sums stays valid.
DgemmView sub_view(const Handle *h, int trans, long begin, long count) {
  DgemmView v = full_view(h, trans);
  if (trans) {
    v.Z += begin * v.pitch;
    v.rows = count;
    v.y = h->d_f + begin;
  } else {
    v.Z += begin / 4;
    v.K = count;
    v.x = h->d_f + begin;
  }
  return v;
}

// C (+)= -2 (x^T B) y^T, the centring term of the view (reference dgemm_compressed_cuda.cu:421-459)
int dgemm_center_init(Handle *h, const DgemmView &v, int n, const double *dB, long ldb, double *dC,
                      long ldc, int centered, double *w, const int *guard) {
  if (centered) {
    MB_CUDA(cudaMemsetAsync(w, 0, sizeof(double) * n, h->stream));
    int ysplit = (int)(ceil_div(v.K, 256 * 16) < 64 ? ceil_div(v.K, 256 * 16) : 64);
    if (ysplit < 1) ysplit = 1;
    center_w_kernel<<<dim3(n, ysplit), 256, 0, h->stream>>>(dB, ldb, v.K, v.x, w, guard);
    count_launch();
    MB_LAUNCH_CHECK();
  }
  return dgemm_init_c(h, v, n, dC, ldc, centered, w, guard);
}

// C (+)= w y^T from centring scalars that are already computed (dgemm_tc.cu col_stats_kernel)
int dgemm_init_c(Handle *h, const DgemmView &v, int n, double *dC, long ldc, int centered,
                 const double *w, const int *guard) {
  if (centered || !v.accumulate) {
    init_c_kernel<<<(unsigned)ceil_div(v.rows, 256), 256, 0, h->stream>>>(
        dC, ldc, v.rows, n, w, v.y, centered, v.accumulate, guard);
    count_launch();
    MB_LAUNCH_CHECK();
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// mode 0: fp64 SIMT streaming kernel.  Block = 8 warps x 4 rows; K walked in chunks of 512
// genotypes (= one 128-byte line per row, read coalesced, one 32-bit word per lane); the B
// chunk is staged in shared memory (stride-17 padding -> conflict-free 64-bit reads).
// Split-K over gridDim.y with fp64 atomics into the centre-initialised C.
// ---------------------------------------------------------------------------------------------
constexpr int F64_ROWS = 32;
constexpr int F64_KCHUNK = 512;
constexpr int F64_BPITCH = F64_KCHUNK + F64_KCHUNK / 16;  // 544

template <int NC>
__global__ void __launch_bounds__(256)
dgemm_fp64_kernel(const uint8_t *__restrict__ A, long pitch, long rows, long K,
                  const double *__restrict__ B, long ldb, int ncols, double *__restrict__ C,
                  long ldc, int chunks_per_split) {
  __shared__ double Bs[NC * F64_BPITCH];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long row0 = (long)blockIdx.x * F64_ROWS + warp * 4;
  const long nchunks = (K + F64_KCHUNK - 1) / F64_KCHUNK;
  const long ch_begin = (long)blockIdx.y * chunks_per_split;
  long ch_end = ch_begin + chunks_per_split;
  if (ch_end > nchunks) ch_end = nchunks;

  double acc[4][NC];
#pragma unroll
  for (int r = 0; r < 4; r++)
#pragma unroll
    for (int c = 0; c < NC; c++) acc[r][c] = 0.0;

  for (long ch = ch_begin; ch < ch_end; ch++) {
    const long kbase = ch * F64_KCHUNK;
    __syncthreads();
    for (int idx = tid; idx < NC * F64_KCHUNK; idx += 256) {
      const int col = idx / F64_KCHUNK, kk = idx % F64_KCHUNK;
      const long k = kbase + kk;
      Bs[col * F64_BPITCH + kk + (kk >> 4)] =
          (k < K && col < ncols) ? B[k + (size_t)col * ldb] : 0.0;
    }
    __syncthreads();
    uint32_t w[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const long row = row0 + r;
      w[r] = row < rows
                 ? __ldg(reinterpret_cast<const uint32_t *>(A + row * pitch + ch * 128) + lane)
                 : 0u;
    }
#pragma unroll
    for (int j = 0; j < 16; j++) {
      double d[4];
#pragma unroll
      for (int r = 0; r < 4; r++) {
        const uint32_t v = (w[r] >> (2 * j)) & 3u;
        d[r] = v ? __hiloint2double((int)(0x3FE00000u + (v << 20)), 0) : 0.0;
      }
#pragma unroll
      for (int c = 0; c < NC; c++) {
        // field j of the word is genotype 4 (j & 3) + (j >> 2) of the 16 (interleaved layout)
        const double b = Bs[c * F64_BPITCH + lane * 17 + 4 * (j & 3) + (j >> 2)];
#pragma unroll
        for (int r = 0; r < 4; r++) acc[r][c] = fma(d[r], b, acc[r][c]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 4; r++)
#pragma unroll
    for (int c = 0; c < NC; c++) {
      double v = acc[r][c];
      for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && row0 + r < rows && c < ncols) atomicAdd(&C[row0 + r + (size_t)c * ldc], v);
    }
}

template <int NC>
static int launch_fp64(Handle *h, const uint8_t *A, long pitch, long rows, long K, const double *B,
                       long ldb, int ncols, double *C, long ldc) {
  const long row_tiles = ceil_div(rows, F64_ROWS);
  const long nchunks = ceil_div(K, F64_KCHUNK);
  long ksplit = ceil_div(4L * h->sm_count, row_tiles);
  if (ksplit > nchunks) ksplit = nchunks;
  if (ksplit < 1) ksplit = 1;
  if (ksplit > 65535) ksplit = 65535;
  const int cps = (int)ceil_div(nchunks, ksplit);
  ksplit = ceil_div(nchunks, cps);
  dgemm_fp64_kernel<NC><<<dim3((unsigned)row_tiles, (unsigned)ksplit), 256, 0, h->stream>>>(
      A, pitch, rows, K, B, ldb, ncols, C, ldc, cps);
  count_launch();
  MB_LAUNCH_CHECK();
  return 0;
}

int dgemm_fp64_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC,
                      long ldc) {
  const uint8_t *A = trans ? h->d_snp_major : h->d_indiv_major;
  const long pitch = trans ? h->pitch_s : h->pitch_i;
  const long rows = trans ? h->snps : h->indiv, K = trans ? h->indiv : h->snps;
  for (int c0 = 0; c0 < n;) {
    const int left = n - c0;
    const double *B = dB + (size_t)c0 * ldb;
    double *C = dC + (size_t)c0 * ldc;
    int used;
    if (left >= 10)     { used = 10; MB_CHECK(launch_fp64<10>(h, A, pitch, rows, K, B, ldb, used, C, ldc)); }
    else if (left > 4)  { used = left < 8 ? left : 8; MB_CHECK(launch_fp64<8>(h, A, pitch, rows, K, B, ldb, used, C, ldc)); }
    else if (left > 2)  { used = left; MB_CHECK(launch_fp64<4>(h, A, pitch, rows, K, B, ldb, used, C, ldc)); }
    else if (left == 2) { used = 2; MB_CHECK(launch_fp64<2>(h, A, pitch, rows, K, B, ldb, used, C, ldc)); }
    else                { used = 1; MB_CHECK(launch_fp64<1>(h, A, pitch, rows, K, B, ldb, used, C, ldc)); }
    c0 += used;
  }
  return 0;
}

int dgemm_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc) {
  if (n <= 0) return 0;
  const Options &o = options();
  if (o.dgemm_mode == 1 && dgemm_tc_available())
    return dgemm_tc_device(h, trans, n, dB, ldb, dC, ldc);
  MB_CHECK(dgemm_center_init(h, full_view(h, trans), n, dB, ldb, dC, ldc, o.centered, h->d_w, nullptr));
  const int ev = (int)(h->ev_count % Handle::kEvRing);
  h->ev_trans[ev] = trans;
  MB_CUDA(cudaEventRecord(h->ev_begin[ev], h->stream));
  MB_CHECK(dgemm_fp64_device(h, trans, n, dB, ldb, dC, ldc));
  MB_CUDA(cudaEventRecord(h->ev_end[ev], h->stream));
  h->ev_count++;
  return 0;
}

}  // namespace mb200
