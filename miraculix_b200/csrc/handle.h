// handle.h -- the opaque object behind `void *compressed` and the process-global options.
// Replaces struct GPU_gemm_storage (reference src/cuda/dgemm_compressed_cuda.h:87-100) and the
// ten flags of setOptions_compressed that reach the GPU path (src/miraculix/options.h:27-57).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>
#include "stager.h"

namespace mb200 {

struct Options {
  int use_gpu = 1;
  int cores = 0;
  int centered = 1;  // genetics.centered: 1 = RowMeans (subtract 2 f), 0 = raw Z
  int print_details = 0;
  int variant = 256;
  int dgemm_mode = 1;      // 0 fp64 SIMT, 1 int8-slice tcgen05
  int dgemm_slices = 0;    // 8-bit digits per column of B in mode 1; 0 = adaptive (6, verified, else 8)
  int crossprod_mode = 2;  // 0 dp4a SIMT, 1 tcgen05 one CTA per tile, 2 tcgen05 CTA pairs (cta_group::2)
  int wide_cluster = 0;    // B wider than one column group: 1 = clusters with multicast packed tiles (dgemm_tc.cu)
  double small_limit = 4.0e7;      // host-pointer dgemm_compressed with snps*indiv*n at most this: one fused fp64 kernel (dgemm_small.cu)
  int pipe_chunks = 4;             // SNP chunks of the host-pointer dgemm_compressed pipeline
  double pipe_min_bytes = 1.0e7;   // ... with at least this many bytes of the snps x n panel per chunk
};
Options &options();

constexpr uint32_t kHandleMagic = 0x4d423230u;  // "MB20"

// HBM layout: both orientations value-coded (2-bit field = genotype 0/1/2, PLINK
// missing -> 0), rows padded with zeros to a multiple of 128 bytes (= 512 genotypes), base
// pointers 256-byte aligned (cudaMalloc).  Inside every little-endian 32-bit word (16 genotypes
// k = 16q .. 16q+15 of a row) the fields are interleaved: genotype 16q + 4j + b lives at bits
// [8b + 2j, 8b + 2j + 1], so (word >> 2j) & 0x03030303 yields genotypes 16q+4j .. 16q+4j+3 as
// four int8 -- the tensor-core kernel expands 16 genotypes with 7 integer instructions.
struct Handle {
  uint32_t magic = kHandleMagic;
  int device = 0;
  long snps = 0, indiv = 0;
  int max_n = 0;
  long pitch_s = 0;  // bytes per SNP row   (snp-major:  [snps][pitch_s],  K = indiv)
  long pitch_i = 0;  // bytes per indiv row (indiv-major:[indiv][pitch_i], K = snps)
  uint8_t *d_snp_major = nullptr;
  uint8_t *d_indiv_major = nullptr;
  bool zeroed = false;  // layouts memset to 0 (only the generic repack kernel needs it, repack.cu)
  double *d_f = nullptr;                   // allele frequencies (caller-supplied), length snps
  unsigned long long *d_counts = nullptr;  // allele counts per SNP from the repack
  std::vector<double> h_f;                 // host copy for get_compressed_freq
  // per-call scratch, sized for max_n columns, allocated once (no per-call malloc):
  double *d_B = nullptr;   // [max(snps,indiv)] x max_n, ld = rows of B
  double *d_C = nullptr;   // [max(snps,indiv)] x max_n
  double *d_w = nullptr;   // centring scalars, column maxima, exponents, flag: 8 (max_n + 16) doubles
  int last_slices = 0;            // digits used by the last tensor-core dgemm on this handle
  bool last_adaptive = false;
  double rowsum_max[2] = {0, 0};  // max_r sum_k z_rk of the indiv-major / snp-major layout
  int8_t *d_Bq = nullptr;  // sliced int8 panel for the tensor-core path
  size_t bq_bytes = 0;
  void *d_cg = nullptr;    // conjugate-gradient workspace (solver.cu), allocated on first use
  size_t cg_bytes = 0;
  void *h_cg = nullptr;    // pinned scalars the solver's kernels write for the host
  Stager stager;           // pinned ring for pageable caller panels (stager.h), created on first use
  cudaStream_t stream = nullptr;
  // host-pointer dgemm_compressed pipelines the PCIe copies of B / C against the kernels:
  cudaStream_t copy_stream = nullptr;
  static constexpr int kPipeMax = 16;
  cudaEvent_t ev_pipe[kPipeMax] = {nullptr};
  void *h_pipe = nullptr;  // pinned word: the precision flag travels back with the result
  // pinned staging of the latency path (dgemm_small.cu): just the two small panels, grown on demand --
  // a first call must not pay for the 24 MB ring (20-40 ms of page pinning)
  uint8_t *h_small = nullptr;
  size_t h_small_bytes = 0;
  // measured rate of this handle's pinned host copies (GB/s; [0] H2D of 'N', [1] D2H of 'T'): events
  // around the largest chunk of the last call; the chunk count of the next call follows from it
  // (boxes whose GPUs share host links deliver a fraction of a link per device)
  cudaEvent_t ev_copy[2] = {nullptr, nullptr};
  double link_gbs[2] = {0, 0};
  int copy_dir = -1;       // direction of the measurement in flight, -1 = none
  double copy_bytes = 0;
  // ring of event pairs around the MAIN kernel of each dgemm_compressed call (bench.py reads
  // the per-launch durations of the timed region through mb200_dgemm_kernel_times)
  static constexpr int kEvRing = 128;
  cudaEvent_t ev_begin[kEvRing] = {nullptr}, ev_end[kEvRing] = {nullptr};
  int ev_trans[kEvRing] = {0};
  unsigned long long ev_count = 0;
  int sm_count = 148;
};

// A sub-problem of one orientation: rows x K genotypes starting at Z (row pitch `pitch` bytes).
// 'N' views cut the SNP range (K) at multiples of 512 SNPs, 'T' views cut the SNP rows.
struct DgemmView {
  const uint8_t *Z = nullptr;
  long pitch = 0, rows = 0, K = 0;
  const double *x = nullptr;  // centring weights along K    (f for 'N', nullptr = ones for 'T')
  const double *y = nullptr;  // centring weights along rows (f for 'T', nullptr = ones for 'N')
  double rowsum_max = 0;      // upper bound of max_r sum_k z_rk over the view
  int trans = 0;
  int accumulate = 0;         // 1: C already holds earlier K ranges, add instead of initialise
};
DgemmView full_view(const Handle *h, int trans);
DgemmView sub_view(const Handle *h, int trans, long begin, long count);  // SNP range [begin, begin+count)

Handle *as_handle(void *p);  // validates magic, sets device; nullptr + status on failure
// device < 0: CUDA_DEVICE / the calling thread's current device (select_device)
int create_handle(long snps, long indiv, int max_n, Handle **out, int device = -1);
void destroy_handle(Handle *h);
int repack_from_host(Handle *h, const uint8_t *plink);
int finish_handle(Handle *h, const double *f_host);  // f (host, or nullptr = from counts) + row-sum maxima

// repack.cu
int repack_from_device(Handle *h, const uint8_t *d_plink);  // d_plink: [snps][ceil(indiv/4)]
// rows [s_offset, s_offset + snps) of the simulated matrix with this seed (shards of one matrix)
int simulate_plink(uint8_t *d_plink, long snps, long indiv, unsigned long long seed,
                   cudaStream_t stream, long s_offset = 0);
int freq_from_counts(Handle *h);  // d_f = counts / (2 indiv)
int rowsum_maxima(Handle *h);     // fills h->rowsum_max (synchronises the stream)

// dgemm_small.cu: latency path of the host-pointer call on small problems
bool dgemm_small_applies(const Handle *h, int n);
int dgemm_small_host(Handle *h, int trans, int n, const double *B, long ldb, double *C, long ldc);
int dgemm_small_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc);

// dgemm.cu
int dgemm_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc);
// SNP-sharded 'N' (multi.cu): the digit count is fixed by the caller for the whole matrix and the
// a-posteriori check needs the REDUCED result, so the shard only reports the rigorous bound of its
// partial: d_err[c] = rowsum_max 2^(e_c - P - 1) (0 for a zero column).
struct TcExternal {
  int S = 0;                // digits per column (3..8)
  double *d_err = nullptr;  // device, n doubles
  bool no_sync = false;     // leave the handle's stream running (the caller orders what follows on it)
};
// Host panels B, C: copies pipelined against the kernels in SNP chunks (dgemm_tc.cu)
int dgemm_tc_host_pipelined(Handle *h, int trans, int n, const double *B, long ldb, double *C,
                            long ldc, int b_dev, int c_dev, const TcExternal *ext = nullptr);
int dgemm_tc_choose_digits(int n, long rows, long K, bool *adaptive);  // the policy of mb200_set_dgemm_mode
double dgemm_tc_tolerance();

int dgemm_tc_read_trace(long long *out, int n);
int dgemm_tc_last_precision(Handle *h, int *slices);  // dgemm_tc.cu, MB200_TRACE=1

// What the epilogue does with a finished accumulator tile.  kind 0: store int32 counts into C32
// (+= when accumulating K slabs).  kind 1/2 (last K slab): convert to fp64 -- adding the int32
// partial of earlier slabs when accumulating -- apply the GRM centring/scaling of
// crossproduct.jl:95-106 on the way out (affine) and write the final result directly:
//   1: full symmetric column-major n x n matrix: every upper entry (i <= j) goes to (i,j) and (j,i)
//   2: compact row-major panel [local rows][n] of one partition's row tiles (crossprod.cu)
struct CpFinish {
  int kind = 0;
  int affine = 0, do_scale = 0;
  int part = 0, nparts = 1;
  long n = 0;
  double *out = nullptr;
  const double *m = nullptr;     // row sums of the count matrix
  const double *scal = nullptr;  // [0] = sum m, [1] = 2 sum f (1 - f)
  int *tail = nullptr;           // scratch tiles of the K-split tail (crossprod_tc.cu), set by the launcher
};

// crossprod.cu
struct CrossprodTiming {
  double gemm_ms = 0, epilogue_ms = 0, executed_ops = 0;
};
int crossprod_device(const uint8_t *d_x, long k, long rows, int is_plink_format, int mode,
                     const double *d_f, int do_scale, double *d_ans, cudaStream_t stream,
                     CrossprodTiming *timing);

long crossprod_partition_rows(long rows, int part, int nparts);
// Work list of a band of row tiles, built and uploaded once (synchronises `stream`: do it before
// the pipeline fills); freed stream-ordered by crossprod_band_plan_free.
struct BandPlan {
  void *d_tiles = nullptr;
  int num_tiles = 0, real_tiles = 0, pair = 0, tile_begin = 0, tile_end = 0;
};
int crossprod_band_plan(long rows, int tile_begin, int tile_end, cudaStream_t stream, BandPlan *plan);
void crossprod_band_plan_free(BandPlan *plan, cudaStream_t stream);
int crossprod_band_counts(const uint8_t *d_x, long k, long rows, int is_plink_format, const BandPlan &plan,
                          int32_t *C32, cudaStream_t stream, double *executed_ops);
int crossprod_row_sums(const uint8_t *d_x, long k, long rows, int is_plink_format, const double *d_f,
                       int do_scale, double *d_m, cudaStream_t stream);
int crossprod_partition_device(const uint8_t *d_x, long k, long rows, int is_plink_format, int mode,
                               const double *d_f, int do_scale, int part, int nparts,
                               double *d_panel, cudaStream_t stream, CrossprodTiming *timing);

}  // namespace mb200
