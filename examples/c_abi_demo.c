/* c_abi_demo.c -- the reference's call sequence (tests/dgemm_compressed/test_5codesapi.f90:120-209,
 * docs/genotype_matrix_multiplication.md:4-46) written in plain C against libmiraculix_b200.so:
 *
 *   setOptions_compressed -> plink2compressed -> dgemm_compressed 'N' / 'T' -> free_compressed
 *
 * and checked against a naive loop over the bed.  Build (from the repo root):
 *   gcc -std=c99 -O2 -Iinclude examples/c_abi_demo.c -o c_abi_demo \
 *       -Lmiraculix_b200 -lmiraculix_b200 -Wl,-rpath,$PWD/miraculix_b200 -lm
 * Needs a B200 to run; without one plink2compressed reports the missing device and leaves the
 * handle NULL, which the program prints and returns 2 (tests/test_abi.py builds and runs it). */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include "miraculix_b200.h"

static int geno(const unsigned char *bed, long bs, long s, long i) { /* PLINK code -> 0 / 1 / 2, missing -> 0 */
  const int code = (bed[s * bs + i / 4] >> (2 * (i % 4))) & 3;
  return code == 3 ? 2 : code == 2 ? 1 : 0;
}

int main(void) {
  const int snps = 1003, indiv = 203, n = 3;
  const long bs = (indiv + 3) / 4;
  unsigned char *bed = calloc((size_t)snps * bs, 1);
  double *f = malloc(sizeof(double) * snps);
  unsigned long long state = 88172645463325252ull;
  for (long s = 0; s < snps; s++) {
    long sum = 0;
    for (long i = 0; i < indiv; i++) {
      state ^= state << 13; state ^= state >> 7; state ^= state << 17;     /* xorshift64 */
      const int g = (int)(state % 3);
      const int code = g == 2 ? 3 : g == 1 ? 2 : 0;
      bed[s * bs + i / 4] |= (unsigned char)(code << (2 * (i % 4)));
      sum += g;
    }
    f[s] = (double)sum / (2.0 * indiv);
  }
  /* use_gpu, cores, floatLoop, meanSubstract, ignore_missings, do_not_center, do_normalize,
   * use_miraculix_freq, variant, print_details: the vector of utils/benchmark/benchmark.f90:186 */
  setOptions_compressed(1, 1, 0, 0, 1, 0, 0, 0, 256, 0);
  void *h = NULL;
  plink2compressed((char *)bed, (char *)bed, snps, indiv, f, n, &h);
  if (!h) {
    printf("no handle: %s\n", mb200_last_message());
    return 2;
  }
  int bad = 0;
  for (int t = 0; t < 2; t++) {
    const long k = t ? indiv : snps, m = t ? snps : indiv;
    double *B = malloc(sizeof(double) * k * n), *C = malloc(sizeof(double) * m * n);
    for (long j = 0; j < k * n; j++) B[j] = sin(0.37 * (double)j);
    dgemm_compressed(t ? "T" : "N", h, n, B, (int)k, C, (int)m);
    double worst = 0.0, scale = 0.0;
    for (int c = 0; c < n; c++)
      for (long r = 0; r < m; r++) {
        double want = 0.0;
        for (long j = 0; j < k; j++) {
          const long s = t ? r : j, i = t ? j : r;
          want += ((double)geno(bed, bs, s, i) - 2.0 * f[s]) * B[j + c * k];
        }
        const double d = fabs(C[r + c * m] - want);
        if (d > worst) worst = d;
        if (fabs(want) > scale) scale = fabs(want);
      }
    printf("dgemm_compressed '%c': max|dC| / max|C| = %.2e\n", t ? 'T' : 'N', worst / scale);
    if (!(worst <= 1e-10 * scale)) bad = 1;
    free(B);
    free(C);
  }
  free_compressed(&h);
  if (h) bad = 1; /* free_compressed nulls the caller's pointer */
  free(bed);
  free(f);
  printf(bad ? "FAILED\n" : "ok\n");
  return bad;
}
