/* oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's algorithm for the compressed-genotype hot path.
 * Nothing under miraculix_b200/ may include, link or call this file; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg do, and only as the checker.
 *
 * All citations are file:line relative to /root/reference.  The fp64 dgemm_compressed
 * semantics restated here are pinned against the UNMODIFIED reference CPU "5codes" library
 * (oracle/_ref/ref_miraculix.so, built by oracle/Makefile) in tests/test_oracle.py; the
 * PLINK->2bit byte table is pinned against the 256 constants of
 * src/cuda/snp_multiply_cuda.h:202 (tests/golden/plink2twobit_lut.npy).
 *
 * Conventions:
 *   Z            genotype matrix indiv x snps with entries 0/1/2
 *   plink        SNP-major .bed payload (no 3-byte header): snps rows of ceil(indiv/4) bytes,
 *                four 2-bit codes per byte, lowest bits = first individual
 *   PLINK code   00 -> 0, 01 (missing) -> 0, 10 -> 1, 11 -> 2
 *                (src/miraculix/MXinfo.h:143 PLINK2HUMAN; src/cuda/dgemm_compressed_cuda.h:259-267)
 *   twobit code  the code is the value (src/bindings/Julia/compressed_operations.jl:95-108)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define EXPORT __attribute__((visibility("default")))

static inline int plink2human(int c) { return c <= 1 ? 0 : c - 1; } /* MXinfo.h:143 */

/* ---------------------------------------------------------------------------------------
 * PLINK -> 2bit byte table.  Restates create_plink_conversion_table
 * (src/bindings/Julia/read_plink.jl:50-80): walk the byte from the most significant pair to
 * the least, shifting the result left by two each time; "10" -> 1, "11" -> 2, "00" -> 0 and
 * any "01" pair makes the WHOLE byte the NaN code 0xff.  Must equal d_conversion_table
 * (src/cuda/snp_multiply_cuda.h:202).
 * ------------------------------------------------------------------------------------- */
EXPORT void oracle_plink2twobit_lut(uint8_t table[256]) {
  for (int v = 0; v < 256; v++) {
    uint8_t out = 0;
    for (int pos = 3; pos >= 0; pos--) {
      int c = (v >> (2 * pos)) & 3;
      out = (uint8_t)(out << 2);
      if (c == 2) out |= 1;
      else if (c == 3) out |= 2;
      else if (c == 1) { out = 0xff; break; }
    }
    table[v] = out;
  }
}

/* ---------------------------------------------------------------------------------------
 * Host transpose of a packed 2-bit matrix.  Restates transpose_genotype_matrix
 * (src/bindings/Julia/compressed_operations.jl:45-66; Fortran twin
 * src/bindings/Fortran/modplink_miraculix.f90:177-218): input  [snps][ceil(indiv/4)],
 * output [indiv][ceil(snps/4)], codes copied verbatim, unused bit pairs zero.
 * ------------------------------------------------------------------------------------- */
EXPORT void oracle_transpose_2bit(const uint8_t *in, long rows, long cols, uint8_t *out) {
  long ib = (cols + 3) / 4, ob = (rows + 3) / 4;
  memset(out, 0, (size_t)cols * ob);
  for (long r = 0; r < rows; r++)
    for (long c = 0; c < cols; c++) {
      int code = (in[r * ib + c / 4] >> (2 * (c % 4))) & 3;
      out[c * ob + r / 4] |= (uint8_t)(code << (2 * (r % 4)));
    }
}

/* ---------------------------------------------------------------------------------------
 * Device layout of the B200 build (DESIGN.md "data layout in HBM"), restated on the host so
 * that the GPU repack can be compared byte for byte:
 *   snp_major  [snps ][pitch_s]  pitch_s = 128*ceil(ceil(indiv/4)/128)
 *   ind_major  [indiv][pitch_i]  pitch_i = 128*ceil(ceil(snps /4)/128)
 * every 2-bit field holds the genotype VALUE 0/1/2 (PLINK2HUMAN applied, missing -> 0),
 * padding bytes and padding bit pairs are zero.  Transposition semantics as above.
 * Within each aligned group of 16 genotypes (one little-endian 32-bit word of the row) the
 * fields are interleaved: genotype k = 16q + 4j + b sits in byte 4q + b at bits [2j, 2j+1],
 * so that (word >> 2j) & 0x03030303 is four consecutive genotypes as four int8.
 * ------------------------------------------------------------------------------------- */
static inline long dev_byte(long k) { return 4 * (k / 16) + (k % 4); }
static inline int dev_shift(long k) { return 2 * (int)((k % 16) / 4); }

EXPORT void oracle_repack(const uint8_t *plink, long snps, long indiv,
                          uint8_t *snp_major, long pitch_s,
                          uint8_t *ind_major, long pitch_i) {
  long bs = (indiv + 3) / 4;
  memset(snp_major, 0, (size_t)snps * pitch_s);
  memset(ind_major, 0, (size_t)indiv * pitch_i);
  for (long s = 0; s < snps; s++)
    for (long i = 0; i < indiv; i++) {
      int v = plink2human((plink[s * bs + i / 4] >> (2 * (i % 4))) & 3);
      snp_major[s * pitch_s + dev_byte(i)] |= (uint8_t)(v << dev_shift(i));
      ind_major[i * pitch_i + dev_byte(s)] |= (uint8_t)(v << dev_shift(s));
    }
}

/* Inverse of the in-row interleave: packed device rows -> one value per byte, plain order. */
EXPORT void oracle_unpack_device(const uint8_t *packed, long rows, long pitch, long cols,
                                 uint8_t *out) {
  for (long r = 0; r < rows; r++)
    for (long k = 0; k < cols; k++)
      out[r * cols + k] = (packed[r * pitch + dev_byte(k)] >> dev_shift(k)) & 3;
}

/* Allele counts per SNP: sum_i z_is (popcount rule of read_plink.jl:197-203 when there are
 * no missings; with missings this is the sum of decoded values). */
EXPORT void oracle_allele_counts(const uint8_t *plink, long snps, long indiv, int64_t *counts) {
  long bs = (indiv + 3) / 4;
  for (long s = 0; s < snps; s++) {
    int64_t acc = 0;
    for (long i = 0; i < indiv; i++)
      acc += plink2human((plink[s * bs + i / 4] >> (2 * (i % 4))) & 3);
    counts[s] = acc;
  }
}

/* ---------------------------------------------------------------------------------------
 * dgemm_compressed.  Restates the reference GPU routine
 * (src/cuda/dgemm_compressed_cuda.cu:218-489) = CPU gV_vG_means_double with external f
 * (src/miraculix/Vector.matrix.D.cc:42-229):
 *   trans == 0 ('N'):  C[indiv x n] = Z   * B[snps  x n] - 1 * (2 f^T B)
 *   trans == 1 ('T'):  C[snps  x n] = Z^T * B[indiv x n] - f * (2 1^T B)
 * centered == 0 drops the rank-1 term.  B, C column-major with leading dimensions ldb, ldc.
 * Accumulates in long double so the oracle is at least as accurate as either implementation.
 * ------------------------------------------------------------------------------------- */
EXPORT void oracle_dgemm_compressed(int trans, const uint8_t *plink, long snps, long indiv,
                                    const double *f, int centered, int n, const double *B,
                                    long ldb, double *C, long ldc) {
  long bs = (indiv + 3) / 4;
  long m = trans ? snps : indiv, k = trans ? indiv : snps;
  for (int col = 0; col < n; col++) {
    const double *b = B + (size_t)col * ldb;
    double *c = C + (size_t)col * ldc;
    long double w = 0.0L; /* f^T b ('N') or 1^T b ('T') */
    for (long j = 0; j < k; j++) w += trans ? (long double)b[j] : (long double)f[j] * b[j];
#pragma omp parallel for schedule(static)
    for (long r = 0; r < m; r++) {
      long double acc = 0.0L;
      for (long j = 0; j < k; j++) {
        long s = trans ? r : j, i = trans ? j : r;
        int v = plink2human((plink[s * bs + i / 4] >> (2 * (i % 4))) & 3);
        if (v) acc += (long double)v * b[j];
      }
      if (centered) acc -= 2.0L * w * (trans ? (long double)f[r] : 1.0L);
      c[r] = (double)acc;
    }
  }
}

/* ---------------------------------------------------------------------------------------
 * Integer crossproduct  M = X X^T  (snp_multiply_gpu, src/cuda/snp_multiply_cuda.cu:38-380).
 * X has `rows` rows (the reference's `indiv` argument) of ceil(k/4) bytes (k = the
 * reference's `snps` argument), 2-bit fields, the field IS the value 0..3 -- exactly what
 * the two masked mma.sync.u4 passes compute (src/cuda/snp_multiply_cuda.h:170-192).  With
 * is_plink_format every byte first goes through the PLINK->2bit table (:225-239, :370-373),
 * so a byte with a missing code becomes 0xff = four 3s, as in the reference.  Result:
 * rows x rows doubles, column-major, symmetric (:330-340).  Exact (int64 accumulate).
 * ------------------------------------------------------------------------------------- */
EXPORT void oracle_crossprod(const uint8_t *x, long k, long rows, double *ans,
                             int is_plink_format) {
  long nb = (k + 3) / 4;
  uint8_t lut[256];
  oracle_plink2twobit_lut(lut);
  int8_t *dec = (int8_t *)malloc((size_t)rows * nb * 4);
#pragma omp parallel for schedule(static)
  for (long r = 0; r < rows; r++)
    for (long b = 0; b < nb; b++) {
      uint8_t v = x[r * nb + b];
      if (is_plink_format) v = lut[v];
      for (int t = 0; t < 4; t++) dec[(r * nb + b) * 4 + t] = (int8_t)((v >> (2 * t)) & 3);
    }
#pragma omp parallel for schedule(dynamic, 8)
  for (long i = 0; i < rows; i++)
    for (long j = i; j < rows; j++) {
      const int8_t *a = dec + i * nb * 4, *b = dec + j * nb * 4;
      int64_t acc = 0;
      for (long t = 0; t < nb * 4; t++) acc += (int)a[t] * (int)b[t];
      ans[i + j * rows] = ans[j + i * rows] = (double)acc;
    }
  free(dec);
}

/* ---------------------------------------------------------------------------------------
 * GRM epilogue.  Restates grm() of src/bindings/Julia/crossproduct.jl:91-106 in the order
 * written there (math: docs/grm.md:3-10):
 *   m = colsums(M); M += (-1/n) m 1^T; M += (-1/n) 1 m^T; M += sum(m)/n^2; if do_scale M /= c
 *   c = 2 sum_j f_j (1 - f_j)
 * M is n x n column-major, updated in place.
 * ------------------------------------------------------------------------------------- */
EXPORT void oracle_grm_epilogue(double *M, long n, const double *f, long snps, int do_scale) {
  double *m = (double *)calloc((size_t)n, sizeof(double));
  for (long j = 0; j < n; j++) {
    double acc = 0.0;
    for (long i = 0; i < n; i++) acc += M[i + j * n];
    m[j] = acc;
  }
  double alpha = -1.0 / (double)n;
  for (long j = 0; j < n; j++)
    for (long i = 0; i < n; i++) M[i + j * n] += alpha * m[i] * 1.0; /* ger!(alpha, m, 1, M) */
  for (long j = 0; j < n; j++)
    for (long i = 0; i < n; i++) M[i + j * n] += alpha * 1.0 * m[j]; /* ger!(alpha, 1, m, M) */
  double tot = 0.0;
  for (long j = 0; j < n; j++) tot += m[j];
  double shift = tot / ((double)n * (double)n);
  for (long t = 0; t < n * n; t++) M[t] += shift;
  if (do_scale) {
    double c = 0.0;
    for (long j = 0; j < snps; j++) c += f[j] * (1.0 - f[j]);
    c *= 2.0;
    for (long t = 0; t < n * n; t++) M[t] /= c;
  }
  free(m);
}

/* ---------------------------------------------------------------------------------------
 * LD epilogue.  Restates ld() of src/bindings/Julia/crossproduct.jl:128-152:
 *   M <- M - 4 n f f^T (upper triangle via syr!, mirrored), sigma = sqrt(diag M),
 *   M <- M ./ sigma ./ sigma^T.   M is snps x snps, n = number of individuals.
 * ------------------------------------------------------------------------------------- */
EXPORT void oracle_ld_epilogue(double *M, long snps, long n_indiv, const double *f) {
  double alpha = -4.0 * (double)n_indiv;
  for (long j = 0; j < snps; j++)
    for (long i = 0; i <= j; i++) {
      double v = M[i + j * snps] + alpha * f[i] * f[j];
      M[i + j * snps] = v;
      M[j + i * snps] = v;
    }
  double *sig = (double *)malloc((size_t)snps * sizeof(double));
  for (long j = 0; j < snps; j++) sig[j] = sqrt(M[j + j * snps]);
  for (long j = 0; j < snps; j++)
    for (long i = 0; i < snps; i++) M[i + j * snps] = M[i + j * snps] / sig[i] / sig[j];
  free(sig);
}
