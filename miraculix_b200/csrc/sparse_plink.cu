// sparse_plink.cu -- sparse_times_plink and dgemm_plink on the GPU.
//
// Replaces, behind the same C ABI (Fortran binding src/bindings/Fortran/mod5codesapi.f90:84-100,
// test tests/sparse_plink/test_sparse_plink.f90:99):
//   sparse_times_plink   src/miraculix/5codesAPI.c:131-157 -> sparseTGenoPlinkApi
//                        (5codesChar.cc:472-492) -> sparseTGeno (haplogeno.cc:1685-1713: memset of
//                        C, BUG on a transposed sparse matrix) -> sparseTGenoPlink
//                        (plinkUint.cc:303-352: per nonzero, PLINK code <= 1 -> 0, 10 -> 1, 11 -> 2)
//   dgemm_plink          5codesAPI.c:112-127 -> vectorGenoPlinkApi (5codesChar.cc:495-519), which
//                        is a BUG() stub in the reference (plink256.cc:332, and f != NULL: BUG);
//                        here it is the one-shot product plink2compressed + dgemm_compressed + free.
// The reference implements both on the CPU only.
#include <stdlib.h>
#include <vector>
#include "../../include/miraculix_b200.h"
#include "common.cuh"
#include "handle.h"

namespace mb200 {

// C[j, s] = sum_{nz of sparse row j} val * z(code[col][s]),  code: R rows x lda bytes of PLINK
// 2-bit codes, W codes per row.  One warp = one sparse row x 512 codes (128 bytes per code row,
// one 32-bit word = 16 codes per lane); a block = 8 sparse rows.  The 8 x 512 results are staged
// in shared memory so the column-major C is written in runs of 8 consecutive doubles.
constexpr int SP_ROWS = 8;
constexpr int SP_CODES = 512;

__global__ void __launch_bounds__(256)
sparse_plink_kernel(const uint8_t *__restrict__ code, long lda, long W, int nIdx,
                    const int *__restrict__ rowIdx, const int *__restrict__ colIdx,
                    const double *__restrict__ val, double *__restrict__ C, long ldc) {
  __shared__ double tile[SP_CODES][SP_ROWS + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long s0 = (long)blockIdx.x * SP_CODES;          // first code (column of C) of the block
  const int j = blockIdx.y * SP_ROWS + warp;             // sparse row of this warp
  double acc[16];
#pragma unroll
  for (int t = 0; t < 16; t++) acc[t] = 0.0;
  const long byte0 = s0 / 4 + 4 * lane;                   // this lane's 4 bytes of a code row
  if (j < nIdx && byte0 < lda) {
    const int e0 = rowIdx[j], e1 = rowIdx[j + 1];
    for (int e = e0; e < e1; e++) {
      const uint8_t *row = code + (long)colIdx[e] * lda + byte0;
      const double v = val[e];
      uint32_t w = 0;                                     // rows need not be word aligned
#pragma unroll
      for (int b = 0; b < 4; b++)
        if (byte0 + b < lda) w |= (uint32_t)row[b] << (8 * b);
#pragma unroll
      for (int t = 0; t < 16; t++) {
        const uint32_t c = (w >> (2 * t)) & 3u;           // 00, 01 (missing) -> 0; 10 -> 1; 11 -> 2
        acc[t] += c == 2u ? v : c == 3u ? 2.0 * v : 0.0;
      }
    }
  }
#pragma unroll
  for (int t = 0; t < 16; t++) tile[16 * lane + t][warp] = acc[t];
  __syncthreads();
  for (int e = threadIdx.x; e < SP_CODES * SP_ROWS; e += 256) {
    const int s = e / SP_ROWS, r = e % SP_ROWS;
    const long gs = s0 + s;
    const int gj = blockIdx.y * SP_ROWS + r;
    if (gs < W && gj < nIdx) C[gj + gs * ldc] = tile[s][r];
  }
}

}  // namespace mb200

using namespace mb200;

static int sp_is_trans(const char *t) {  // 5codesAPI.c:73-77
  if (*t == 'T' || *t == 't' || *t == 'Y' || *t == 'y') return 1;
  if (*t != 'N' && *t != 'n') exit(99);
  return 0;
}

extern "C" {

void sparse_times_plink(char *transsparse, char *transcompressed, char *plink, char *plink_transposed,
                        int snps, int indiv, int nIdx, int *rowIdxB, int *colIdxB, double *B,
                        double *C, int Ldc) {
  clear_status();
  if (sp_is_trans(transsparse)) {  // haplogeno.cc:1702: `if (tSparse) BUG;`
    fprintf(stderr, "miraculix_b200: sparse_times_plink: a transposed sparse matrix is not "
                    "supported (the reference aborts here as well)\n");
    exit(EXIT_FAILURE);
  }
  // 'N': rows of the code matrix = SNPs (plink), C is nIdx x indiv, sparse columns index SNPs;
  // 'T': rows = individuals (plink_transposed), C is nIdx x snps  (5codesAPI.c:146-154)
  const int tc = sp_is_trans(transcompressed);
  const uint8_t *code = reinterpret_cast<const uint8_t *>(tc ? plink_transposed : plink);
  const long R = tc ? indiv : snps, W = tc ? snps : indiv, lda = (W + 3) / 4;
  if (!code || !rowIdxB || !C || nIdx < 0 || Ldc < nIdx || R <= 0 || W <= 0) {
    fail(-2, __func__, __LINE__, "bad arguments (nIdx=%d Ldc=%d rows=%ld codes=%ld)", nIdx, Ldc, R, W);
    return;
  }
  if (select_device() < 0) return;
  const long nnz = nIdx > 0 ? rowIdxB[nIdx] : 0;
  uint8_t *d_code = nullptr;
  int *d_row = nullptr, *d_col = nullptr;
  double *d_val = nullptr, *d_C = nullptr;
  int rc = 0;
  cudaError_t e;
#define SP_CUDA(expr)                                                                    \
  if (!rc && (e = (expr)) != cudaSuccess)                                                 \
    rc = fail((int)e, __func__, __LINE__, "%s: %s", #expr, cudaGetErrorString(e))
  SP_CUDA(cudaMalloc(&d_code, (size_t)R * lda));
  SP_CUDA(cudaMalloc(&d_row, sizeof(int) * (nIdx + 1)));
  SP_CUDA(cudaMalloc(&d_col, sizeof(int) * (nnz + 1)));
  SP_CUDA(cudaMalloc(&d_val, sizeof(double) * (nnz + 1)));
  SP_CUDA(cudaMalloc(&d_C, sizeof(double) * (size_t)Ldc * W));
  SP_CUDA(cudaMemcpy(d_code, code, (size_t)R * lda, cudaMemcpyHostToDevice));
  SP_CUDA(cudaMemcpy(d_row, rowIdxB, sizeof(int) * (nIdx + 1), cudaMemcpyHostToDevice));
  if (nnz > 0) {
    SP_CUDA(cudaMemcpy(d_col, colIdxB, sizeof(int) * nnz, cudaMemcpyHostToDevice));
    SP_CUDA(cudaMemcpy(d_val, B, sizeof(double) * nnz, cudaMemcpyHostToDevice));
  }
  SP_CUDA(cudaMemset(d_C, 0, sizeof(double) * (size_t)Ldc * W));  // haplogeno.cc:1700: whole Ldc x W
  if (!rc && nIdx > 0) {
    // column indices must address rows of the code matrix
    for (long i = 0; i < nnz && !rc; i++)
      if (colIdxB[i] < 0 || colIdxB[i] >= R)
        rc = fail(-2, __func__, __LINE__, "colIdxB[%ld] = %d outside [0, %ld)", i, colIdxB[i], R);
    if (!rc) {
      dim3 grid((unsigned)ceil_div(W, SP_CODES), (unsigned)ceil_div(nIdx, SP_ROWS));
      sparse_plink_kernel<<<grid, 256>>>(d_code, lda, W, nIdx, d_row, d_col, d_val, d_C, Ldc);
      count_launch();
      SP_CUDA(cudaGetLastError());
    }
  }
  SP_CUDA(cudaMemcpy(C, d_C, sizeof(double) * (size_t)Ldc * W, cudaMemcpyDeviceToHost));
#undef SP_CUDA
  cudaFree(d_code);
  cudaFree(d_row);
  cudaFree(d_col);
  cudaFree(d_val);
  cudaFree(d_C);
}

void dgemm_plink(char *trans, char *plink, char *plink_transposed, int snps, int indiv, double *f,
                 int n, double *B, int Ldb, double *C, int Ldc) {
  clear_status();
  const int t = sp_is_trans(trans);
  // one-shot: handle -> product -> free.  f == NULL (the only call the reference's stub accepts
  // before it BUGs) = raw Z product; with f the product is centred like dgemm_compressed.
  std::vector<double> zero;
  const Options saved = options();
  if (!f) {
    zero.assign((size_t)snps, 0.0);
    options().centered = 0;
  }
  void *h = nullptr;
  plink2compressed(plink, plink_transposed, snps, indiv, f ? f : zero.data(), n, &h);
  if (h) {
    char tr[2] = {t ? 'T' : 'N', 0};
    dgemm_compressed(tr, h, n, B, Ldb, C, Ldc);
    const int rc = status().code;
    free_compressed(&h);
    if (rc) status().code = rc;
  }
  options() = saved;
}

}  // extern "C"
