// dgemm_small.cu -- dgemm_compressed on small problems: ONE kernel per call.
//
// BASELINE configs[0] (the reference's own test, tests/dgemm_compressed: 10k SNPs x 1k individuals,
// one column) moves 2.6 MB: its time is the host's API calls, not the GPU.  The tensor-core path
// costs about fifteen of them per call (memsets, column statistics, quantisation, tensor maps, main
// kernel, a-posteriori check); the reference costs more (per-call cublasCreate, mallocs, Dgeam,
// Dgemv, n Daxpy: dgemm_compressed_cuda.cu:218-489).  Below a work limit the host-pointer call
// therefore takes this path: upload B, one launch, download C, one synchronisation.
//
// The kernel fuses the centring (dgemm_compressed_cuda.cu:421-459) into the product:
//   C[r,c] = sum_k z_rk B[k,c]  -  2 y_r sum_k x_k B[k,c]        (x, y) = (f, 1) for 'N', (1, f) for 'T'
// A warp owns R consecutive rows and a range of 512-genotype chunks (one 128-byte line per row:
// lane l holds word l of each).  B is read COALESCED -- lane l takes genotype 32 j + l of the chunk,
// j = 0..15, once for the R rows -- and the 2-bit field that goes with it comes from lane
// 2 j + (l >> 4) by a shuffle, at a field position that is fixed per lane (handle.h: genotype 4 a + b
// of a word sits at bits 8 b + 2 a).  No shared-memory staging, no block-wide barrier in the loop,
// no atomics: the K ranges of one row sit in ONE block and are summed in a fixed order, so results
// repeat bit for bit; plain fp64 FMAs (no digit slices) keep them at fp64 rounding.
#include "common.cuh"
#include "handle.h"

namespace mb200 {

constexpr int SK_KCHUNK = 512;

// Block = 8 warps = (8 / KS) row groups x KS ranges of K.  NC = columns per launch, ncols <= NC real.
template <int NC, int KS, int R>
__global__ void __launch_bounds__(256)
dgemm_small_kernel(const uint8_t *__restrict__ A, long pitch, long rows, long K,
                   const double *__restrict__ B, long ldb, int ncols, const double *__restrict__ x,
                   const double *__restrict__ y, int centered, double *__restrict__ C, long ldc) {
  constexpr int RG = 8 / KS;      // row groups per block
  constexpr int JB = NC <= 2 ? 16 / NC : NC == 4 ? 4 : 2;  // j per batch of loads (16 doubles of B in flight)
  constexpr int NV = R * NC + NC; // values a warp hands over: az[r][c], then aw[c]
  __shared__ double red[8][NV];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ks = warp % KS;
  const long row0 = ((long)blockIdx.x * RG + warp / KS) * R;
  const long nchunks = (K + SK_KCHUNK - 1) / SK_KCHUNK;
  const int g = lane & 15;                                 // my genotype inside its word, every j
  const int shift = 2 * (4 * (g & 3) + (g >> 2));          // ... and where its field sits
  const int src0 = lane >> 4;

  double az[R][NC], aw[NC];
#pragma unroll
  for (int c = 0; c < NC; c++) {
    aw[c] = 0.0;
#pragma unroll
    for (int r = 0; r < R; r++) az[r][c] = 0.0;
  }

  if (row0 < rows) {
    const uint8_t *a0 = A + row0 * pitch + 4 * lane;
    uint32_t cur[R], nxt[R];
#pragma unroll
    for (int r = 0; r < R; r++)
      cur[r] = (ks < nchunks && row0 + r < rows) ? __ldg(reinterpret_cast<const uint32_t *>(a0 + r * pitch + (long)ks * 128)) : 0u;
    for (long ch = ks; ch < nchunks; ch += KS) {
      const long chn = ch + KS;  // the next chunk's words are in flight while this one is consumed
#pragma unroll
      for (int r = 0; r < R; r++)
        nxt[r] = (chn < nchunks && row0 + r < rows) ? __ldg(reinterpret_cast<const uint32_t *>(a0 + r * pitch + chn * 128)) : 0u;
      const long kbase = ch * SK_KCHUNK + lane;
      const bool full = (ch + 1) * SK_KCHUNK <= K;
      // all loads of a batch of j are issued before the first one is used: a warp that walks its
      // two or three chunks alone must not pay one L2 round trip per j
#pragma unroll
      for (int j0 = 0; j0 < 16; j0 += JB) {
        double b[JB][NC], xk[JB];
#pragma unroll
        for (int jj = 0; jj < JB; jj++) {
          const long k = kbase + 32 * (j0 + jj);
          const bool ok = full || k < K;
#pragma unroll
          for (int c = 0; c < NC; c++) b[jj][c] = (ok && c < ncols) ? B[k + (size_t)c * ldb] : 0.0;
          xk[jj] = (centered && ok) ? (x ? x[k] : 1.0) : 0.0;
        }
#pragma unroll
        for (int jj = 0; jj < JB; jj++) {
          const int j = j0 + jj;
#pragma unroll
          for (int c = 0; c < NC; c++) aw[c] = fma(xk[jj], b[jj][c], aw[c]);
#pragma unroll
          for (int r = 0; r < R; r++) {
            const uint32_t wj = __shfl_sync(0xffffffffu, cur[r], 2 * j + src0);
            const uint32_t v = (wj >> shift) & 3u;
            const double d = v ? __hiloint2double((int)(0x3FE00000u + (v << 20)), 0) : 0.0;  // 1 -> 1.0, 2 -> 2.0
#pragma unroll
            for (int c = 0; c < NC; c++) az[r][c] = fma(d, b[jj][c], az[r][c]);
          }
        }
      }
#pragma unroll
      for (int r = 0; r < R; r++) cur[r] = nxt[r];
    }
  }
#pragma unroll
  for (int c = 0; c < NC; c++) {
#pragma unroll
    for (int r = 0; r < R; r++) {
      for (int o = 16; o; o >>= 1) az[r][c] += __shfl_xor_sync(0xffffffffu, az[r][c], o);
      if (lane == 0) red[warp][r * NC + c] = az[r][c];
    }
    for (int o = 16; o; o >>= 1) aw[c] += __shfl_xor_sync(0xffffffffu, aw[c], o);
    if (lane == 0) red[warp][R * NC + c] = aw[c];
  }
  __syncthreads();
  if (ks == 0) {  // one thread per (row, column) sums the K ranges in a fixed order
    for (int idx = lane; idx < R * NC; idx += 32) {
      const int r = idx / NC, c = idx % NC;
      const long row = row0 + r;
      if (row < rows && c < ncols) {
        double z = 0.0, w = 0.0;
#pragma unroll
        for (int i = 0; i < KS; i++) {
          z += red[warp + i][idx];
          w += red[warp + i][R * NC + c];
        }
        C[row + (size_t)c * ldc] = centered ? fma(-2.0 * (y ? y[row] : 1.0), w, z) : z;
      }
    }
  }
}

template <int NC>
static int launch_small(Handle *h, const DgemmView &v, const double *B, long ldb, int ncols, double *C,
                        long ldc, int centered) {
  constexpr int R = NC <= 2 ? 8 : 4;
  // split K over the warps of a block while the row groups alone cannot fill the device
  const long nchunks = ceil_div(v.K, SK_KCHUNK), groups = ceil_div(v.rows, R);
  int ks = 1;
  while (ks < 8 && 2 * ks <= nchunks && ceil_div(groups * ks, 8) < 2L * h->sm_count) ks *= 2;
#define SK_LAUNCH(KS)                                                                              \
  dgemm_small_kernel<NC, KS, R><<<(unsigned)ceil_div(groups, 8 / KS), 256, 0, h->stream>>>(          \
      v.Z, v.pitch, v.rows, v.K, B, ldb, ncols, v.x, v.y, centered, C, ldc)
  if (ks == 8) SK_LAUNCH(8);
  else if (ks == 4) SK_LAUNCH(4);
  else if (ks == 2) SK_LAUNCH(2);
  else SK_LAUNCH(1);
#undef SK_LAUNCH
  count_launch();
  MB_LAUNCH_CHECK();
  return 0;
}

// device panels, asynchronous on the handle's stream
int dgemm_small_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc) {
  const DgemmView v = full_view(h, trans);
  const int centered = options().centered;
  for (int c0 = 0; c0 < n;) {
    const int left = n - c0;
    const double *B = dB + (size_t)c0 * ldb;
    double *C = dC + (size_t)c0 * ldc;
    int used;
    if (left >= 8)      { used = 8; MB_CHECK(launch_small<8>(h, v, B, ldb, used, C, ldc, centered)); }
    else if (left > 2)  { used = left < 4 ? left : 4; MB_CHECK(launch_small<4>(h, v, B, ldb, used, C, ldc, centered)); }
    else if (left == 2) { used = 2; MB_CHECK(launch_small<2>(h, v, B, ldb, used, C, ldc, centered)); }
    else                { used = 1; MB_CHECK(launch_small<1>(h, v, B, ldb, used, C, ldc, centered)); }
    c0 += used;
  }
  return 0;
}

// Does the host-pointer call take this path?  Work limit in genotype x column products
// (mb200_set_dgemm_small_limit / MB200_SMALL_LIMIT, 0 = never); panels of at most 4 MB each.
bool dgemm_small_applies(const Handle *h, int n) {
  const Options &o = options();
  if (o.small_limit <= 0 || o.dgemm_mode != 1 || o.dgemm_slices != 0) return false;
  const double work = (double)h->snps * (double)h->indiv * (double)n;
  const long big = h->snps > h->indiv ? h->snps : h->indiv;
  return work <= o.small_limit && (double)big * n * 8.0 <= 4.0e6;
}

// Host panels: upload, one launch per column group, download, ONE synchronisation.  Pageable
// panels (what the bindings pass) are copied by the host into / out of a small pinned buffer of the
// handle: cudaMemcpyAsync on pageable memory synchronises inside the driver, twice per call.
int dgemm_small_host(Handle *h, int trans, int n, const double *B, long ldb, double *C, long ldc) {
  const long m = trans ? h->snps : h->indiv, k = trans ? h->indiv : h->snps;
  const bool b_pin = host_is_pinned(B), c_pin = host_is_pinned(C);
  const double *srcB = B;
  long ld_src = ldb;
  double *dstC = C;
  long ld_dst = ldc;
  if (!b_pin || !c_pin) {
    const size_t b_bytes = b_pin ? 0 : round_up((long)sizeof(double) * k * n, 256);
    const size_t c_bytes = c_pin ? 0 : sizeof(double) * m * n;
    if (b_bytes + c_bytes > h->h_small_bytes) {
      if (h->h_small) MB_CUDA(cudaFreeHost(h->h_small));
      h->h_small = nullptr;
      h->h_small_bytes = 0;
      const size_t want = (size_t)round_up((long)(b_bytes + c_bytes), 65536);
      MB_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&h->h_small), want, cudaHostAllocPortable));
      h->h_small_bytes = want;
    }
    if (!b_pin) {
      double *slot = reinterpret_cast<double *>(h->h_small);
      parallel_copy_2d(slot, sizeof(double) * k, B, sizeof(double) * ldb, sizeof(double) * k, (size_t)n);
      srcB = slot;
      ld_src = k;
    }
    if (!c_pin) {
      dstC = reinterpret_cast<double *>(h->h_small + b_bytes);
      ld_dst = m;
    }
  }
  if (ld_src == k)
    MB_CUDA(cudaMemcpyAsync(h->d_B, srcB, sizeof(double) * k * n, cudaMemcpyHostToDevice, h->stream));
  else
    MB_CUDA(cudaMemcpy2DAsync(h->d_B, sizeof(double) * k, srcB, sizeof(double) * ld_src, sizeof(double) * k,
                              (size_t)n, cudaMemcpyHostToDevice, h->stream));
  const int ev = (int)(h->ev_count % Handle::kEvRing);
  h->ev_trans[ev] = trans;
  MB_CUDA(cudaEventRecord(h->ev_begin[ev], h->stream));
  MB_CHECK(dgemm_small_device(h, trans, n, h->d_B, k, h->d_C, m));
  MB_CUDA(cudaEventRecord(h->ev_end[ev], h->stream));
  h->ev_count++;
  h->last_slices = 0;  // fp64 arithmetic, no digit slices
  h->last_adaptive = false;
  if (ld_dst == m)
    MB_CUDA(cudaMemcpyAsync(dstC, h->d_C, sizeof(double) * m * n, cudaMemcpyDeviceToHost, h->stream));
  else
    MB_CUDA(cudaMemcpy2DAsync(dstC, sizeof(double) * ld_dst, h->d_C, sizeof(double) * m, sizeof(double) * m,
                              (size_t)n, cudaMemcpyDeviceToHost, h->stream));
  MB_CUDA(cudaStreamSynchronize(h->stream));
  if (!c_pin)
    parallel_copy_2d(C, sizeof(double) * ldc, dstC, sizeof(double) * m, sizeof(double) * m, (size_t)n);
  return 0;
}

}  // namespace mb200
