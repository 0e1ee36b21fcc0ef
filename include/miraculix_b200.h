/* miraculix_b200.h -- C ABI of the B200-native compressed-genotype library.
 *
 * PART 1 re-exports, symbol for symbol, what the reference's bindings bind for this path
 * (citations relative to the reference tree):
 *   - Fortran bind(C) interfaces   src/bindings/Fortran/mod5codesapi.f90:22-82
 *   - Julia ccall sites            src/bindings/Julia/dgemm_compressed.jl:55-57,88,129-131,154
 *                                  src/bindings/Julia/crossproduct.jl:54-58
 *   - C declarations               src/miraculix/5codes.h:137-153, src/cuda/snp_multiply_cuda.h:113-114
 * so that `libmiraculix_b200.so` can be dlopen'ed in place of `src/miraculix/miraculix.so`
 * (see INTEGRATION.md).  PART 2 is additive (prefix mb200_): device-pointer variants so an
 * iterative solver can stay on the GPU, the fused GRM / LD epilogues, introspection used by the
 * parity tests, and an on-device simulator for benchmark-sized inputs.
 *
 * All arrays in PART 1 are caller-owned HOST memory borrowed for the duration of the call.
 * There is no CPU fallback anywhere: every entry point runs hand-written sm_100a kernels and
 * reports an error (stderr + status) if no CUDA device is usable.
 */
#ifndef MIRACULIX_B200_H
#define MIRACULIX_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * PART 1 -- reference-compatible surface
 * ---------------------------------------------------------------------------------------- */

/* replaces src/miraculix/5codesAPI.c:43-70 (-> setOptions5, 5codesChar.cc:165-209).
 * Process-global.  The GPU-compatible combination is enforced exactly like the reference
 * (5codesChar.cc:192-193): use_miraculix_freq must be 0, ignore_missings must be 1,
 * do_normalize must be 0, otherwise: message on stderr and exit(EXIT_FAILURE).
 * use_gpu == 0 is accepted but the work still runs on the GPU (this build has no CPU path). */
void setOptions_compressed(int use_gpu, int cores, int floatLoop, int meanSubstract,
                           int ignore_missings, int do_not_center, int do_normalize,
                           int use_miraculix_freq, int variant, int print_details);

/* replaces src/miraculix/5codesAPI.c:80-96 (-> plinkToSxI 5codesChar.cc:344-399 -> plink2gpu
 * src/cuda/dgemm_compressed_cuda.cu:43-170).
 * plink: SNP-major .bed payload, snps rows of ceil(indiv/4) bytes.  plink_transposed is
 * accepted for signature compatibility and NOT read (may be NULL or alias plink): both device
 * orientations are produced by the GPU repack.  f: allele frequencies, length snps (copied).
 * On failure *compressed is set to NULL (Julia checks this, dgemm_compressed.jl:90). */
void plink2compressed(char *plink, char *plink_transposed, int snps, int indiv, double *f,
                      int max_n, void **compressed);

/* replaces src/miraculix/5codesAPI.c:98-110 (-> dgemm_compressed_gpu,
 * src/cuda/dgemm_compressed_cuda.cu:218-489).
 *   'N'/'n':          C[indiv x n] = (Z - 2 1 f^T)   B[snps  x n]
 *   'T'/'t'/'Y'/'y':  C[snps  x n] = (Z - 2 1 f^T)^T B[indiv x n]
 * any other first byte: exit(99) as 5codesAPI.c:73-77.  B, C column-major; Ldb/Ldc are
 * honoured (the reference GPU path silently assumes Ldb = rows(B), Ldc = rows(C)).
 * n > max_n is an error (message, C untouched) instead of the reference's scratch overflow. */
void dgemm_compressed(char *trans, void *compressed, int n, double *B, int Ldb, double *C, int Ldc);

/* replaces src/miraculix/5codesAPI.c:159-161 (-> freegpu dgemm_compressed_cuda.cu:176-213).
 * Unlike the reference this also sets *compressed = NULL. */
void free_compressed(void **compressed);

/* replaces src/miraculix/5codesAPI.c:37-39; works for GPU handles (host copy of f kept). */
void get_compressed_freq(void *compressed, double *f);

/* replaces src/cuda/snp_multiply_cuda.cu:375-382 (-> snp_crossproduct :38-368).
 * snp_matrix: `indiv` rows of ceil(snps/4) bytes, 2-bit fields; the field is the value unless
 * is_plink_format, in which case every byte first goes through the reference's PLINK->2bit
 * table (src/cuda/snp_multiply_cuda.h:202).  ans: indiv x indiv doubles, column-major, full
 * symmetric matrix of exact integer counts.  Returns 0 on success, 1 on error. */
int snp_multiply_gpu(unsigned char *snp_matrix, int snps, int indiv, double *ans,
                     bool is_plink_format);

/* replaces src/cuda/solve_cuda.cu:947-951 (-> dense_solve :69-278), declared in
 * src/cuda/solve_cuda.h:84-86 and bound by src/bindings/Julia/solve.jl:172-173; it follows the GRM
 * in examples/gblup/calculate_gblup.jl.  Solves A X = B for symmetric positive definite A
 * (input_size x input_size, only the lower triangle of the column-major view is read), B and X
 * input_size x rhs_cols column-major; *logdet (optional) = log det A.  *status = 0 / 1.
 * cuSOLVER potrf/potrs underneath (library call). */
void potrs_solve_gpu(double *A, unsigned int input_size, double *B, unsigned int rhs_cols, double *X,
                     double *logdet, int oversubscribe, int *status);

/* replaces src/miraculix/5codesAPI.c:131-157 (-> sparseTGenoPlinkApi 5codesChar.cc:472-492 ->
 * sparseTGeno haplogeno.cc:1685-1713 -> sparseTGenoPlink plinkUint.cc:303-352; CPU only in the
 * reference), bound by src/bindings/Fortran/mod5codesapi.f90:84-100.  CSR sparse matrix
 * (rowIdxB[nIdx+1], colIdxB, B; 0-based) times the PLINK-coded genotypes, uncentred, missing = 0:
 *   transcompressed 'N': C[nIdx x indiv] = S[nIdx x snps]  Z^T   (rows of `plink` = SNPs)
 *   transcompressed 'T': C[nIdx x snps ] = S[nIdx x indiv] Z     (rows of `plink_transposed`)
 * C column-major with leading dimension Ldc; the whole Ldc x columns block is zeroed first, as
 * the reference does.  transsparse must be 'N' (the reference aborts otherwise; so do we). */
void sparse_times_plink(char *transsparse, char *transcompressed, char *plink, char *plink_transposed,
                        int snps, int indiv, int nIdx, int *rowIdxB, int *colIdxB, double *B,
                        double *C, int Ldc);
/* replaces src/miraculix/5codesAPI.c:112-127.  In the reference this entry is a BUG() stub
 * (vectorGenoPlinkApi 5codesChar.cc:495-519 -> plink256.cc:332); here it is the one-shot product
 * plink2compressed + dgemm_compressed + free_compressed: f == NULL gives the raw Z product, f !=
 * NULL the centred one.  plink_transposed is not read. */
void dgemm_plink(char *trans, char *plink, char *plink_transposed, int snps, int indiv, double *f,
                 int n, double *B, int Ldb, double *C, int Ldc);

/* ------------------------------------------------------------------------------------------
 * Several GPUs behind the SAME calls (the reference is single-device and selects it
 * with CUDA_DEVICE, src/cuda/cuda_utils.cu:188-248).  Environment, read when a handle is created
 * or a crossproduct is called:
 *   MIRACULIX_NGPUS=N     shard over devices CUDA_DEVICE .. CUDA_DEVICE+N-1 (CUDA_DEVICE default 0)
 *   CUDA_DEVICE=0,2,5     shard over an explicit list
 *   TILE_SIZE=rows        crossproduct: cap on the row height of the int32 accumulation panel a
 *                         device holds per pass (the reference's memory knob,
 *                         src/cuda/snp_multiply_cuda.cu:68-72); MB200_GRM_ACC_GB caps its bytes
 *   PRINT_LEVEL           >= 0: banner + one line per call, > 0: per-call details (cuda_utils.cu:44-82)
 * plink2compressed then returns ONE handle for the whole matrix, sharded by 512-SNP blocks: one
 * host thread per device inside the library, 'T' all-gathers the replicated B over NVLink (every
 * device uploads 1/N of it), 'N' reduce-scatters the partial C (every device downloads 1/N), the
 * CG exchanges indiv doubles per product -- all with the library's own peer-memory kernels, one
 * process, no NCCL.  snp_multiply_gpu / mb200_grm split the output row tiles over the devices and
 * stream panels straight into `ans`.  Unchanged single-process callers (Julia ccall, Fortran
 * bind(C)) therefore use all GPUs.  Caller memory may be pageable: it is staged through pinned
 * rings inside the library, so the copy/compute overlap does not depend on page-locked buffers.
 * ---------------------------------------------------------------------------------------- */

/* ------------------------------------------------------------------------------------------
 * PART 2 -- additive entry points (not in the reference)
 * ---------------------------------------------------------------------------------------- */

/* Programmatic form of MIRACULIX_NGPUS / CUDA_DEVICE=a,b,c for the handles and crossproducts
 * created afterwards; n <= 0 returns to the environment.  A device may be listed more than once
 * (its shards then share the GPU -- exercises the multi-device path on a one-GPU box). */
int mb200_set_devices(int n, const int *devices);
/* Number of device shards behind a handle (1 for an ordinary handle), and shard g as an ordinary
 * single-device handle (owned by the parent; valid for mb200_get_layout / mb200_download_layout /
 * the device-pointer calls) with its SNP range [snp_begin, snp_begin + snp_count). */
int mb200_shard_count(void *compressed);
int mb200_get_shard(void *compressed, int g, void **shard, long *snp_begin, long *snp_count);
/* Handle whose genotypes come from the on-device simulator below (same law and seed rule: the
 * matrix depends on (seed, snps, indiv) only, not on the number of devices); f = allele counts /
 * (2 indiv).  For benchmarks and scale tests: the bed never exists on the host. */
int mb200_simulate_compressed(long snps, long indiv, unsigned long long seed, int max_n,
                              void **compressed);

/* Status of the last call on this thread: 0 ok, else cudaError / internal code; message text. */
int mb200_last_status(void);
const char *mb200_last_message(void);
/* Number of kernels this library has launched since load (bench.py "gpu_launches"). */
unsigned long long mb200_kernel_launches(void);
/* 0: fp64 SIMT streaming kernel.  1: exact int8-slice tcgen05 kernel (default).
 * slices: number of 8-bit digits each column of B is split into: 3..8 fixed, 0 (default) =
 * adaptive: 6 digits, checked after the product against the rigorous bound
 *   rowsum_max * 2^(e_c - 47) <= 5e-11 * max_r |C[r,c]|   for every column c
 * and recomputed with 8 digits (error <= 2^-62 relative to max|B_c|) when the bound fails. */
int mb200_set_dgemm_mode(int mode, int slices);
/* B wider than one group of columns (13 at 6 digits, 10 at 8): 0 (default) = the groups' work units
 * are interleaved in one launch and share the packed tiles in L2; 1 = thread-block clusters of 2 / 4
 * CTAs with the packed tiles multicast (DRAM sees the matrix once; slower in time on B200 because
 * clusters of 4 occupy 132 of 148 SMs -- DESIGN.md 3.2).  Results are identical. */
int mb200_set_dgemm_wide_cluster(int on);
/* Host-pointer dgemm_compressed overlaps the PCIe copies of the snps x n panel (B for 'N', C for
 * 'T') with the kernels by cutting the SNP axis into `chunks` pieces (default 4, at most 16;
 * boundaries multiples of 512 SNPs), fewer when a chunk would hold less than min_bytes of that
 * panel (default 1e7: a 10 MB panel is not cut at all; 0 = always `chunks`; negative = keep).
 * With page-locked panels the handle measures the rate of its own copies (events around the
 * largest chunk of the previous call) and picks the count <= `chunks` that minimises exposed copy
 * + 50 us per chunk: boxes whose GPUs share host links deliver a fraction of a link per device.
 * Results do not depend on the chunking beyond fp64 summation order. */
int mb200_set_pipeline(int chunks, double min_bytes);
/* Host-pointer dgemm_compressed on SMALL problems (snps * indiv * n <= products, panels below
 * 4 MB; default 4e7, env MB200_SMALL_LIMIT; 0 = never) with the default precision policy
 * (mb200_set_dgemm_mode(1, 0)): one fused fp64 kernel -- centring and product -- between one
 * upload and one download, because such calls are bound by the number of API calls, not by the
 * GPU (BASELINE configs[0], the reference's tests/dgemm_compressed case).  fp64 FMAs in a fixed
 * order: reproducible bit for bit; mb200_dgemm_last_precision then reports 0 digits. */
int mb200_set_dgemm_small_limit(double products);
/* 0: SIMT dp4a crossproduct, 1: tcgen05 int8 crossproduct, one CTA per 128 x 256 tile,
 * 2 (default): tcgen05 cta_group::2, CTA pairs on 256 x 256 tiles (full-matrix products and the
 * partitioned GRM, whose pairs take two consecutive row tiles of the partition). */
int mb200_set_crossprod_mode(int mode);

/* dgemm_compressed with DEVICE pointers dB, dC on the handle's device; asynchronous on the
 * handle's stream (mb200_stream).  trans: 0 = 'N', 1 = 'T'.  Returns 0 / error code. */
int mb200_dgemm_compressed_device(int trans, void *compressed, int n, const double *dB, long ldb,
                                  double *dC, long ldc);
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The fp64 dgemm_compressed on DEVICE pointers dB, dC, asynchronous on the handle's
 * stream like mb200_dgemm_compressed_device, but this call allocates nothing: the centring
 * scalars, which the ordinary path keeps in scratch owned by the handle, live in a static
 * __device__ array of MB200_NOALLOC_MAX_COLS = 1024 doubles (8 KiB on every GPU that loads the
 * library, until exit).  n > 1024 is refused with status -2 and C untouched (n is not bound by
 * the handle's max_n).  The array is shared by ALL calls and cleared at the start of each one:
 * calls that overlap on one GPU (other streams, other handles) race on it -- order them.
 * Single-device handles only; honours setOptions_compressed's centring switch, always computes
 * with the fp64 kernel (mb200_set_dgemm_mode does not apply). */
#define MB200_NOALLOC_MAX_COLS 1024
int mb200_dgemm_compressed_noalloc(int trans, void *compressed, int n, const double *dB, long ldb,
                                   double *dC, long ldc);
int mb200_noalloc_max_cols(void); /* = MB200_NOALLOC_MAX_COLS of the loaded library */
/* Same product with each panel on the host OR on the device (flags): host panels are copied over
 * PCIe, pipelined against the kernels like dgemm_compressed.  Blocking (returns after the
 * handle's stream has drained) unless both panels are on the device.  Used by SNP-sharded callers:
 * 'N' with B on the host and C on the device (the partial is all-reduced before it goes home). */
int mb200_dgemm_compressed_mixed(int trans, void *compressed, int n, const double *B, long ldb,
                                 int b_on_device, double *C, long ldc, int c_on_device);
void *mb200_stream(void *compressed); /* cudaStream_t */
/* Device time of the MAIN kernel of the last dgemm_compressed call on this handle (CUDA events
 * recorded on the handle's stream around that one launch); synchronises the stream. */
int mb200_dgemm_last_kernel_ms(void *compressed, float *ms);
/* Durations (ms) of the main kernel of the `last` most recent dgemm_compressed calls on this
 * handle, oldest first (at most 128 are kept); trans[j] (optional) = direction of call j.
 * Synchronises the stream.  Returns the number of entries written, or a negative status. */
int mb200_dgemm_kernel_times(void *compressed, int last, float *ms, int *trans);
/* Precision of the last tensor-core dgemm_compressed on this handle: returns 0 = fixed digit
 * count, 1 = adaptive and the 6-digit product passed its bound, 2 = adaptive and redone with 8
 * digits; *slices = digits of the result.  Synchronises the stream. */
int mb200_dgemm_last_precision(void *compressed, int *slices);
int mb200_sync(void *compressed);
/* Developer aid: with MB200_TRACE=1 the tensor-core dgemm kernel accumulates clock64() segments
 * of one expander warp ([0..7]) and of the MMA warp ([16..19]) of CTA 0; this reads them. */
int mb200_debug_read_trace(long long *out, int n);
/* Developer aid: the host copy pool that stages pageable caller memory (csrc/stager.h), callable
 * without a GPU: dst[r + c*dpitch] = src[r + c*spitch] for r < width bytes, c < height.  Any
 * number of threads may call it at the same time (they share the MB200_COPY_THREADS helpers). */
int mb200_debug_parallel_copy(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                              size_t height);

/* Build a handle from a bed payload that already lives on the device (same layout as the
 * host one).  d_f may be NULL: then f is computed on the device from the allele counts. */
int mb200_plink2compressed_device(const unsigned char *d_plink, long snps, long indiv,
                                  const double *d_f, int max_n, void **compressed);

/* Introspection for the parity tests: device layout of a handle (DESIGN.md "HBM layout"). */
typedef struct {
  long snps, indiv;
  long pitch_snp_major;    /* bytes per SNP row        (multiple of 128) */
  long pitch_indiv_major;  /* bytes per individual row (multiple of 128) */
  const unsigned char *d_snp_major;
  const unsigned char *d_indiv_major;
  const double *d_f;
  const unsigned long long *d_allele_counts; /* sum_i z_is per SNP, computed by the repack */
  int device;
} mb200_layout;
int mb200_get_layout(void *compressed, mb200_layout *out);
/* Download both repacked orientations (padded rows) and the allele counts to host buffers. */
int mb200_download_layout(void *compressed, unsigned char *snp_major, unsigned char *indiv_major,
                          unsigned long long *allele_counts);

/* On-device simulator of PLINK genotypes (benchmark-sized inputs never cross PCIe):
 * p_j ~ U(0.1,0.6), z_ij ~ Binomial(2,p_j), counter-based RNG keyed by (seed, snp, indiv).
 * d_plink: snps x ceil(indiv/4) bytes on the current device. */
int mb200_simulate_plink_device(unsigned char *d_plink, long snps, long indiv,
                                unsigned long long seed);
/* SNP rows [snp_offset, snp_offset + snps) of the matrix with that seed (a shard of it). */
int mb200_simulate_plink_rows_device(unsigned char *d_plink, long snps, long indiv,
                                     unsigned long long seed, long snp_offset);

/* Crossproduct on device-resident input.  d_x: rows x ceil(k/4) bytes.  d_counts (optional,
 * may be NULL): rows x rows int32 row-major upper-triangle-valid scratch is internal; d_ans:
 * rows x rows doubles column-major.  mode: 0 raw counts, 1 centred GRM (crossproduct.jl:91-106;
 * d_f = allele freq length k, do_scale), 2 LD (crossproduct.jl:128-152; n_obs = k). */
int mb200_crossprod_device(const unsigned char *d_x, long k, long rows, int is_plink_format,
                           int mode, const double *d_f, int do_scale, double *d_ans);
/* Host-pointer convenience wrappers of the above (same arguments as snp_multiply_gpu + f). */
int mb200_grm(unsigned char *snp_matrix, int snps, int indiv, double *ans, bool is_plink_format,
              const double *f, bool do_scale);
int mb200_ld(unsigned char *snp_matrix, int snps, int indiv, double *ans, bool is_plink_format,
             const double *f);
/* 2-D output-tile partition of the crossproduct over GPUs (no communication):
 * partition `part` of `nparts` owns the 128-row tiles t with t % nparts == part.  The panel is
 * compact and row-major, [mb200_crossprod_partition_rows(rows, part, nparts)][rows] doubles; local
 * row lr is global row i = ((lr/128) nparts + part) 128 + lr%128; entries with column
 * j < 256 floor(i/256) (strictly left of the diagonal tile) are not computed and read 0 -- use
 * symmetry.  mode 0: raw counts; mode 1: centred (and scaled) GRM, crossproduct.jl:91-106, with
 * the row sums obtained exactly from the input (m = X X^T 1), so ranks never exchange data.
 * d_x: the FULL packed matrix [rows][ceil(k/4)], replicated on every GPU. */
long mb200_crossprod_partition_rows(long rows, int part, int nparts);
int mb200_crossprod_partition_device(const unsigned char *d_x, long k, long rows, int is_plink_format,
                                     int mode, const double *d_f, int do_scale, int part, int nparts,
                                     double *d_panel);
int mb200_grm_partition(unsigned char *snp_matrix, int snps, int indiv, bool is_plink_format, int mode,
                        const double *f, bool do_scale, int part, int nparts, double *panel);
/* Timing of the device part of the last crossproduct call on this thread, in milliseconds:
 * [0] expand+GEMM, [1] epilogue (int32 -> fp64, symmetrise, centre), [2] executed int8 ops. */
int mb200_crossprod_last_timing(double out[3]);

/* potrs_solve_gpu on device-resident data, in place: d_A becomes its Cholesky factor, d_B the
 * solution (so the GRM of mb200_crossprod_device can be solved without leaving the GPU). */
int mb200_potrs_solve_device(double *d_A, long n, double *d_B, long nrhs, double *d_logdet);

/* ------------------------------------------------------------------------------------------
 * PLINK .bed ingest.  Replaces read_bed (src/bindings/Julia/read_plink.jl:
 * 161-228) + transpose_genotype_matrix (compressed_operations.jl:45-66) + plink2compressed: the
 * file `<stem>.bed` (with `<stem>.bim` / `<stem>.fam`, whose line counts give snps / indiv) is
 * checked like read_bed (suffix, companions, magic 6c 1b 01, exact size), streamed through pinned
 * buffers and repacked on the GPU chunk by chunk; the allele frequencies are read_bed's
 * calc_freq rule, set bits per SNP / (2 indiv), and can be fetched with get_compressed_freq.
 * Error codes: -20 not .bed, -21 missing .bim/.fam, -23 bad magic, -25 too small, -26 too large. */
int mb200_bed_shape(const char *bed_path, long *snps, long *indiv); /* validation only, no GPU */
int mb200_read_bed(const char *bed_path, int max_n, void **compressed, long *snps, long *indiv);

/* ------------------------------------------------------------------------------------------
 * Device-resident conjugate gradients (BASELINE configs[3]).
 * Replaces the host loop of examples/iterative_solver/grm_solve_cg.jl:109-133 (GRM_vec :75-85):
 * solves (Zc Zc^T + lambda I) x = b, Zc = Z - 2 1 f^T (or raw Z with do_not_center), with the
 * reference script's recurrences; x holds the start vector on entry and the iterate on return.
 * Stops when ||r|| < tol (checked before every iteration, as :118-121) or after max_iter
 * iterations; *iters = iterations done, *resid_norm = ||r||.  Packed matrix and vectors never
 * leave the GPU; the host reads one scalar per iteration.
 * SNP-sharded solves (one handle per GPU holding the SNP range S_g): pass an
 * `allreduce` hook; it is called once per matrix-vector product with the rank's partial
 * Z[:,S_g] Z[:,S_g]^T p (indiv doubles, device memory) and must sum it in place over all ranks,
 * asynchronously on `stream` (e.g. ncclAllReduce / torch.distributed.all_reduce).  Return 0 on
 * success.  NULL = single GPU. */
typedef int (*mb200_allreduce_fn)(double *d_buf, long count, void *stream, void *user);
int mb200_cg_solve_device(void *compressed, const double *d_b, double *d_x, double lambda,
                          double tol, int max_iter, mb200_allreduce_fn allreduce, void *user,
                          int *iters, double *resid_norm);
/* Same with HOST vectors b, x of length indiv (single GPU). */
int mb200_cg_solve(void *compressed, const double *b, double *x, double lambda, double tol,
                   int max_iter, int *iters, double *resid_norm);

#ifdef __cplusplus
}
#endif
#endif /* MIRACULIX_B200_H */
