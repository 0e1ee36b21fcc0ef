/* TEST INFRASTRUCTURE ONLY (see oracle/Makefile).
 *
 * The reference CPU library embeds a copy of RandomFieldsUtils whose dense solver references
 * nine LAPACK/BLAS symbols.  None of them is reachable from the compressed-genotype path
 * (setOptions_compressed / plink2compressed / dgemm_compressed / free_compressed); they are
 * only needed so that ref_miraculix.so links without a LAPACK in this image.  Reaching one
 * is a bug in the harness, so each stub aborts loudly.
 */
#include <stdio.h>
#include <stdlib.h>

#define STUB(name)                                                              \
  void name(void) {                                                             \
    fprintf(stderr, "oracle/lapack_stubs.c: unexpected call to " #name "\n");   \
    abort();                                                                    \
  }

STUB(dgesdd_)
STUB(dgesv_)
STUB(dgetrf_)
STUB(dgetri_)
STUB(dgetrs_)
STUB(dpotrf_)
STUB(dsyevr_)
STUB(dsyrk_)
STUB(dtrmv_)
