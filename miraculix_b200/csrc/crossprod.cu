// crossprod.cu -- driver of the crossproduct / GRM / LD path on device-resident input
// (SURVEY.md 8a rows a9-a13).  Replaces snp_crossproduct (reference
// src/cuda/snp_multiply_cuda.cu:38-368: per-tile-pair H2D, serialised CUTLASS launches, D2H of
// int32 tiles, host cast + scatter) and the host BLAS epilogues of grm()/ld()
// (src/bindings/Julia/crossproduct.jl:91-106, :128-152).
//
// Pipeline per call, all on one stream:
//   for each K-slab:  expand_2bit_kernel (2-bit -> int8)  ->  GEMM  C32 (+)= X8 X8^T  (upper tiles)
//   counts_to_double_kernel (int32 -> fp64, mirror)  ->  optional GRM / LD epilogue
#include <stdlib.h>
#include "common.cuh"
#include "crossprod_kernels.cuh"
#include "handle.h"

namespace mb200 {

struct TileCoord;
int build_tile_list(long rows, int part, int nparts, TileCoord **d_tiles, int *num_tiles,
                    cudaStream_t stream, int split_kblocks, int grid_max, int *num_tail,
                    int *real_tiles, int pair);
int crossprod_tc_launch(const int8_t *X8, long rows_pad, long kcols, long pitch,
                        const TileCoord *d_tiles,
                        int num_tiles, int32_t *C32, long ld32, int accumulate, int sm_count,
                        cudaStream_t stream, const CpFinish *finish, int num_tail, int pair);
int crossprod_tc_units(int sm_count, int pair);

// Stream-ordered scratch and timing events of one call: released on every exit path (an early
// error return must not leak device memory of the retained pool).
struct CallScratch {
  cudaStream_t stream;
  void *ptrs[8];
  int n = 0;
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  explicit CallScratch(cudaStream_t s) : stream(s) {}
  template <typename T> cudaError_t alloc(T **p, size_t bytes) {
    void *q = nullptr;
    const cudaError_t e = n < 8 ? pool_alloc(&q, bytes, stream) : cudaErrorMemoryAllocation;
    if (e == cudaSuccess) ptrs[n++] = q;
    *p = static_cast<T *>(q);
    return e;
  }
  void adopt(void *p) { if (p && n < 8) ptrs[n++] = p; }
  cudaError_t events() {
    for (int i = 0; i < 3; i++) {
      const cudaError_t e = cudaEventCreate(&ev[i]);
      if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
  }
  ~CallScratch() {
    for (int i = 0; i < n; i++) cudaFreeAsync(ptrs[i], stream);
    for (int i = 0; i < 3; i++) if (ev[i]) cudaEventDestroy(ev[i]);
  }
};

// Side stream + fork/join events of the GRM row sums, created once per host thread and device:
// creating them per call left the GPU idle for ~80 us between the call's first two events.
struct SideStream {
  int device = -1;
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
static int side_stream(int device, SideStream **out) {
  static thread_local SideStream ss;
  if (ss.device != device) {
    if (ss.stream) {
      cudaStreamDestroy(ss.stream);
      cudaEventDestroy(ss.fork);
      cudaEventDestroy(ss.join);
      ss = SideStream();
    }
    MB_CUDA(cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking));
    MB_CUDA(cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming));
    MB_CUDA(cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming));
    ss.device = device;
  }
  *out = &ss;
  return 0;
}

// GRM row sums m = M 1 (d_m[0..rows)), their total (d_m[rows]) and 2 sum f (1 - f) (d_m[rows + 1]),
// exactly from the input: m = X (X^T 1) in 64-bit integers.  overlap: they run on the side stream,
// forked from `stream` here (d_m, d_s allocated on it; the caller's inputs ready); the caller makes
// `stream` wait on *join before the first kernel that reads d_m (the last slab's GEMM epilogue).
static int grm_row_sums(const uint8_t *d_x, long rows, long nb, long k, int is_plink_format,
                        const double *d_f, int do_scale, unsigned int *d_s, double *d_m, int device,
                        cudaStream_t stream, bool overlap, cudaEvent_t *join) {
  cudaStream_t rs = stream;
  SideStream *ss = nullptr;
  if (overlap) {
    MB_CHECK(side_stream(device, &ss));
    rs = ss->stream;
    MB_CUDA(cudaEventRecord(ss->fork, stream));
    MB_CUDA(cudaStreamWaitEvent(rs, ss->fork, 0));
  }
  MB_CHECK(input_row_sums(d_x, rows, nb, is_plink_format, d_s, d_m, rs));
  reduce_sum_kernel<<<1, 1024, 0, rs>>>(d_m, rows, 0, d_m + rows);
  count_launch();
  if (do_scale) {
    reduce_sum_kernel<<<1, 1024, 0, rs>>>(d_f, k, 1, d_m + rows + 1);
    count_launch();
  }
  MB_LAUNCH_CHECK();
  if (overlap) {
    MB_CUDA(cudaEventRecord(ss->join, rs));
    *join = ss->join;
  }
  return 0;
}

static int upload_lut() {
  uint8_t table[256];
  host_plink2twobit(table);
  MB_CUDA(cudaMemcpyToSymbol(c_plink2twobit, table, 256));
  return 0;
}

int crossprod_device(const uint8_t *d_x, long k, long rows, int is_plink_format, int mode,
                     const double *d_f, int do_scale, double *d_ans, cudaStream_t stream,
                     CrossprodTiming *timing) {
  if (rows <= 0 || k <= 0) return fail(-2, __func__, __LINE__, "empty matrix (%ld x %ld)", rows, k);
  // the fused GRM update is exact in 64-bit integers: n^2 M with M <= 4 k must stay below 2^62
  if (mode == 1 && (double)rows * (double)rows * 4.0 * (double)k >= 4.0e18)
    return fail(-2, __func__, __LINE__, "GRM of %ld x %ld exceeds the exact integer epilogue", rows, k);
  int device = 0;
  MB_CUDA(cudaGetDevice(&device));
  const int sms = device_sm_count(device);
  const long nb = ceil_div(k, 4);
  const long rows_pad = round_up(rows, 256);

  // K-slab size: int8 slab of at most MB200_SLAB_GB (default 8) GiB, multiple of 128 columns
  double slab_gb = 8.0;
  if (const char *e = getenv("MB200_SLAB_GB")) slab_gb = atof(e) > 0 ? atof(e) : slab_gb;
  long slab_kbytes = (long)(slab_gb * 1073741824.0 / (double)rows_pad) / 4 / 32 * 32;
  if (slab_kbytes < 32) slab_kbytes = 32;
  if (slab_kbytes > round_up(nb, 32)) slab_kbytes = round_up(nb, 32);
  const long kpad = slab_kbytes * 4;

  if (is_plink_format) MB_CHECK(upload_lut());

  int8_t *X8 = nullptr;
  int32_t *C32 = nullptr;
  double *d_m = nullptr;
  unsigned int *d_s = nullptr;
  TileCoord *d_tiles = nullptr;
  int num_tiles = 0;
  const int tc = options().crossprod_mode >= 1;
  // mode 2: 256 x 256 tiles on CTA pairs (tcgen05 cta_group::2); falls back to one CTA per tile
  // when the device cannot co-schedule pairs
  const int pair = options().crossprod_mode == 2 && crossprod_tc_units(sms, 1) > 0;
  const int nslabs = (int)ceil_div(nb, slab_kbytes);
  // tensor-core path: the last slab's GEMM epilogue writes the final fp64 matrix itself (counts,
  // or the centred/scaled GRM), so the int32 matrix only exists when K needs several slabs
  const bool need_c32 = !tc || nslabs > 1;
  CallScratch scratch(stream);
  MB_CUDA(scratch.events());
  const cudaEvent_t e0 = scratch.ev[0], e1 = scratch.ev[1], e2 = scratch.ev[2];
  MB_CUDA(scratch.alloc(&X8, (size_t)rows_pad * kpad));
  if (need_c32) MB_CUDA(scratch.alloc(&C32, (size_t)rows_pad * rows_pad * sizeof(int32_t)));
  MB_CUDA(scratch.alloc(&d_m, sizeof(double) * (rows + 8)));
  int real_tiles = 0;
  int num_tail = 0;  // single slab: the last partial round of tiles is cut along K (crossprod_tc.cu)
  const int split_kb = nslabs == 1 && !getenv("MB200_NO_TAIL_SPLIT") ? (int)(round_up(nb * 4, 128) / 128) : 0;
  if (tc) {
    MB_CHECK(build_tile_list(rows, 0, 1, &d_tiles, &num_tiles, stream, split_kb,
                             crossprod_tc_units(sms, pair), &num_tail, &real_tiles, pair));
    scratch.adopt(d_tiles);
  }

  MB_CUDA(cudaEventRecord(e0, stream));
  // GRM: the row sums m = M 1 come exactly from the input, m = X (X^T 1) (64-bit integers), so
  // the centring can be applied while the counts leave the accumulators.  They are only needed
  // by the epilogue of the LAST slab's GEMM, so they run on a side stream next to the expansion
  // (both are small HBM/latency-bound kernels) and join right before that launch.
  cudaEvent_t ev_join = nullptr;
  // (single slab only: next to the GEMMs of earlier slabs they cost more than they hide -- measured
  // at configs[4], +8..18 ms on 400)
  const bool overlap = mode == 1 && tc && nslabs == 1 && !getenv("MB200_NO_ROWSUM_OVERLAP");
  if (mode == 1) {
    if (do_scale && !d_f) return fail(-3, __func__, __LINE__, "allele frequencies required for scaling");
    MB_CUDA(scratch.alloc(&d_s, sizeof(unsigned int) * input_row_sums_scratch(nb)));
    MB_CHECK(grm_row_sums(d_x, rows, nb, k, is_plink_format, d_f, do_scale, d_s, d_m, device, stream,
                          overlap, &ev_join));
  }
  // [e0,e1) = row sums when they run in line (SIMT path), ~0 when overlapped; [e1,e2) = the rest
  MB_CUDA(cudaEventRecord(e1, stream));
  CpFinish fin;
  fin.kind = 1;
  fin.affine = mode == 1;
  fin.do_scale = do_scale;
  fin.n = rows;
  fin.out = d_ans;
  fin.m = d_m;
  fin.scal = d_m + rows;
  double executed_ops = 0.0;
  int slab = 0;
  for (long kb0 = 0; kb0 < nb; kb0 += slab_kbytes, slab++) {
    const long kbytes = nb - kb0 < slab_kbytes ? nb - kb0 : slab_kbytes;
    const long kuse = round_up(kbytes * 4, 128);  // columns the GEMM walks this slab
    // expand zero-fills the whole pitch beyond this slab's bytes; the GEMM walks [0, kuse)
    expand_2bit_launch(d_x, rows, nb, kb0, kbytes, is_plink_format, X8, rows_pad, kpad, sms, stream);
    count_launch();
    MB_LAUNCH_CHECK();
    if (tc) {
      if (overlap && slab == nslabs - 1) MB_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
      MB_CHECK(crossprod_tc_launch(X8, rows_pad, kuse, kpad, d_tiles, num_tiles, C32, rows_pad, slab > 0,
                                   sms, stream, slab == nslabs - 1 ? &fin : nullptr, num_tail, pair));
      executed_ops += 2.0 * 128.0 * 256.0 * (double)kuse * (double)real_tiles;
    } else {
      const unsigned g = (unsigned)(rows_pad / 64);
      crossprod_simt_kernel<<<dim3(g, g), 256, 0, stream>>>(X8, kpad, C32, rows_pad, slab > 0);
      count_launch();
      MB_LAUNCH_CHECK();
      executed_ops += 2.0 * 64.0 * 64.0 * (double)kpad * (double)g * (g + 1) / 2.0;
    }
  }
  if (!tc) {  // SIMT cross-check path: separate conversion pass
    const unsigned tb = (unsigned)ceil_div(rows, 32);
    counts_to_double_kernel<<<dim3(tb, tb), 256, 0, stream>>>(C32, rows_pad, rows, d_ans, mode == 1,
                                                              d_m, d_m + rows, do_scale);
    count_launch();
    MB_LAUNCH_CHECK();
  }
  if (mode == 2) {  // LD: rows = SNPs, k = individuals
    if (!d_f) return fail(-3, __func__, __LINE__, "allele frequencies required for LD");
    ld_rank1_kernel<<<sms * 8, 256, 0, stream>>>(d_ans, rows, d_f, -4.0 * (double)k, d_m);
    ld_scale_kernel<<<sms * 8, 256, 0, stream>>>(d_ans, rows, d_m);
    count_launch(2);
    MB_LAUNCH_CHECK();
  }
  MB_CUDA(cudaEventRecord(e2, stream));

  MB_CUDA(cudaStreamSynchronize(stream));  // scratch goes back to the pool when `scratch` leaves scope
  if (timing) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, e0, e1);
    cudaEventElapsedTime(&b, e1, e2);
    timing->gemm_ms = b;      // expand + GEMM with the fused fp64 epilogue (+ LD passes)
    timing->epilogue_ms = a;  // row sums from the input (GRM only)
    timing->executed_ops = executed_ops;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// 2-D output-tile partition over GPUs (BASELINE configs[4]): partition `part` of
// `nparts` owns the 128-row tiles mi with mi % nparts == part (cyclic = balanced over the upper
// triangle) and computes them against all column tiles at or right of the diagonal.  The packed
// input is replicated; nothing is exchanged: even the row sums the centring needs are computed
// exactly from the input, m = X (X^T 1).  Output: compact fp64 panel [local_rows][rows].
// ---------------------------------------------------------------------------------------------
long crossprod_partition_rows(long rows, int part, int nparts) {
  const long Mt = ceil_div(rows, 128);
  const long owned = part < Mt ? ceil_div(Mt - part, nparts) : 0;
  return owned * 128;
}

int crossprod_partition_device(const uint8_t *d_x, long k, long rows, int is_plink_format, int mode,
                               const double *d_f, int do_scale, int part, int nparts,
                               double *d_panel, cudaStream_t stream, CrossprodTiming *timing) {
  if (rows <= 0 || k <= 0) return fail(-2, __func__, __LINE__, "empty matrix (%ld x %ld)", rows, k);
  if (nparts < 1 || part < 0 || part >= nparts)
    return fail(-2, __func__, __LINE__, "bad partition %d of %d", part, nparts);
  if (mode < 0 || mode > 1) return fail(-2, __func__, __LINE__, "mode %d not supported on panels", mode);
  // the fused GRM update is exact in 64-bit integers: n^2 M with M <= 4 k must stay below 2^62
  if (mode == 1 && (double)rows * (double)rows * 4.0 * (double)k >= 4.0e18)
    return fail(-2, __func__, __LINE__, "GRM of %ld x %ld exceeds the exact integer epilogue", rows, k);
  int device = 0;
  MB_CUDA(cudaGetDevice(&device));
  const int sms = device_sm_count(device);
  const long nb = ceil_div(k, 4);
  const long rows_pad = round_up(rows, 256);
  const long local_rows = crossprod_partition_rows(rows, part, nparts);
  if (local_rows == 0) return 0;

  double slab_gb = 8.0;
  if (const char *e = getenv("MB200_SLAB_GB")) slab_gb = atof(e) > 0 ? atof(e) : slab_gb;
  long slab_kbytes = (long)(slab_gb * 1073741824.0 / (double)rows_pad) / 4 / 32 * 32;
  if (slab_kbytes < 32) slab_kbytes = 32;
  if (slab_kbytes > round_up(nb, 32)) slab_kbytes = round_up(nb, 32);
  const long kpad = slab_kbytes * 4;
  if (is_plink_format || mode) MB_CHECK(upload_lut());

  int8_t *X8 = nullptr;
  int32_t *C32 = nullptr;
  double *d_m = nullptr;
  unsigned int *d_s = nullptr;
  TileCoord *d_tiles = nullptr;
  int num_tiles = 0;
  CallScratch scratch(stream);
  MB_CUDA(scratch.events());
  const cudaEvent_t e0 = scratch.ev[0], e1 = scratch.ev[1], e2 = scratch.ev[2];
  MB_CUDA(scratch.alloc(&X8, (size_t)rows_pad * kpad));
  const int nslabs = (int)ceil_div(nb, slab_kbytes);
  // mode 2: CTA pairs over this partition's consecutive row tiles (mi, mi + nparts)
  const int pair = options().crossprod_mode == 2 && crossprod_tc_units(sms, 1) > 0 ? nparts : 0;
  if (nslabs > 1)  // int32 partials only exist when K needs several slabs (an odd tile count's
                   // last pair has a phantom second half: rows padded to whole pairs)
    MB_CUDA(scratch.alloc(&C32, (size_t)round_up(local_rows, 256) * rows_pad * sizeof(int32_t)));
  MB_CUDA(scratch.alloc(&d_m, sizeof(double) * (rows + 8)));
  int real_tiles = 0;
  int num_tail = 0;
  const int split_kb = nslabs == 1 && !getenv("MB200_NO_TAIL_SPLIT") ? (int)(round_up(nb * 4, 128) / 128) : 0;
  MB_CHECK(build_tile_list(rows, part, nparts, &d_tiles, &num_tiles, stream, split_kb,
                           crossprod_tc_units(sms, pair), &num_tail, &real_tiles, pair));
  scratch.adopt(d_tiles);

  MB_CUDA(cudaEventRecord(e0, stream));
  cudaEvent_t ev_join = nullptr;
  const bool overlap = mode == 1 && nslabs == 1 && !getenv("MB200_NO_ROWSUM_OVERLAP");  // see crossprod_device
  if (mode == 1) {
    if (do_scale && !d_f) return fail(-3, __func__, __LINE__, "allele frequencies required for scaling");
    MB_CUDA(scratch.alloc(&d_s, sizeof(unsigned int) * input_row_sums_scratch(nb)));
    MB_CHECK(grm_row_sums(d_x, rows, nb, k, is_plink_format, d_f, do_scale, d_s, d_m, device, stream,
                          overlap, &ev_join));
  }
  // entries left of a row tile's first computed column tile are not part of the panel: zero
  MB_CUDA(cudaMemsetAsync(d_panel, 0, sizeof(double) * (size_t)local_rows * rows, stream));
  MB_CUDA(cudaEventRecord(e1, stream));  // [e0,e1) = panel clear (+ in-line row sums), [e1,e2) = the rest
  CpFinish fin;
  fin.kind = 2;
  fin.affine = mode == 1;
  fin.do_scale = do_scale;
  fin.part = part;
  fin.nparts = nparts;
  fin.n = rows;
  fin.out = d_panel;
  fin.m = d_m;
  fin.scal = d_m + rows;
  double executed_ops = 0.0;
  int slab = 0;
  for (long kb0 = 0; kb0 < nb; kb0 += slab_kbytes, slab++) {
    const long kbytes = nb - kb0 < slab_kbytes ? nb - kb0 : slab_kbytes;
    const long kuse = round_up(kbytes * 4, 128);
    expand_2bit_launch(d_x, rows, nb, kb0, kbytes, is_plink_format, X8, rows_pad, kpad, sms, stream);
    count_launch();
    MB_LAUNCH_CHECK();
    if (overlap && slab == nslabs - 1) MB_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
    MB_CHECK(crossprod_tc_launch(X8, rows_pad, kuse, kpad, d_tiles, num_tiles, C32, rows_pad,
                                 slab > 0, sms, stream, slab == nslabs - 1 ? &fin : nullptr, num_tail, pair));
    executed_ops += 2.0 * 128.0 * 256.0 * (double)kuse * (double)real_tiles;
  }
  MB_CUDA(cudaEventRecord(e2, stream));

  MB_CUDA(cudaStreamSynchronize(stream));
  if (timing) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, e0, e1);
    cudaEventElapsedTime(&b, e1, e2);
    timing->gemm_ms = b;
    timing->epilogue_ms = a;
    timing->executed_ops = executed_ops;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Building blocks of the out-of-core / multi-device host path (crossprod_host.cu): the int32
// counts of a BAND of consecutive row tiles [tile_begin, tile_end) over all K slabs (no fp64
// epilogue: the finishing kernel there converts chunk by chunk while earlier chunks are on their
// way to the host), and the GRM row sums.  C32: [round_up(band rows, 256)][round_up(rows, 256)]
// int32; local row lr is global row 128 tile_begin + lr, valid from column 256 floor(i / 256) on.
// ---------------------------------------------------------------------------------------------
int build_band_tile_list(long rows, int tile_begin, int tile_end, TileCoord **d_tiles, int *num_tiles,
                         cudaStream_t stream, int *real_tiles, int pair);

int crossprod_band_plan(long rows, int tile_begin, int tile_end, cudaStream_t stream, BandPlan *plan) {
  if (tile_begin < 0 || tile_end <= tile_begin || (tile_begin & 1))
    return fail(-2, __func__, __LINE__, "bad band [%d, %d)", tile_begin, tile_end);
  if (options().crossprod_mode < 1)
    return fail(-2, __func__, __LINE__, "the tiled host path needs the tensor-core crossproduct (mode 1 or 2)");
  int device = 0;
  MB_CUDA(cudaGetDevice(&device));
  const int sms = device_sm_count(device);
  plan->pair = options().crossprod_mode == 2 && crossprod_tc_units(sms, 1) > 0 ? 1 : 0;
  plan->tile_begin = tile_begin;
  plan->tile_end = tile_end;
  TileCoord *d_tiles = nullptr;
  MB_CHECK(build_band_tile_list(rows, tile_begin, tile_end, &d_tiles, &plan->num_tiles, stream,
                                &plan->real_tiles, plan->pair));
  plan->d_tiles = d_tiles;
  return 0;
}
void crossprod_band_plan_free(BandPlan *plan, cudaStream_t stream) {
  if (plan->d_tiles) cudaFreeAsync(plan->d_tiles, stream);
  plan->d_tiles = nullptr;
}

int crossprod_band_counts(const uint8_t *d_x, long k, long rows, int is_plink_format, const BandPlan &plan,
                          int32_t *C32, cudaStream_t stream, double *executed_ops) {
  if (rows <= 0 || k <= 0) return fail(-2, __func__, __LINE__, "empty matrix (%ld x %ld)", rows, k);
  int device = 0;
  MB_CUDA(cudaGetDevice(&device));
  const int sms = device_sm_count(device);
  const long nb = ceil_div(k, 4);
  const long rows_pad = round_up(rows, 256);
  double slab_gb = 8.0;
  if (const char *e = getenv("MB200_SLAB_GB")) slab_gb = atof(e) > 0 ? atof(e) : slab_gb;
  long slab_kbytes = (long)(slab_gb * 1073741824.0 / (double)rows_pad) / 4 / 32 * 32;
  if (slab_kbytes < 32) slab_kbytes = 32;
  if (slab_kbytes > round_up(nb, 32)) slab_kbytes = round_up(nb, 32);
  const long kpad = slab_kbytes * 4;
  if (is_plink_format) MB_CHECK(upload_lut());
  int8_t *X8 = nullptr;
  CallScratch scratch(stream);
  MB_CUDA(scratch.alloc(&X8, (size_t)rows_pad * kpad));
  const TileCoord *d_tiles = static_cast<const TileCoord *>(plan.d_tiles);
  int slab = 0;
  for (long kb0 = 0; kb0 < nb; kb0 += slab_kbytes, slab++) {
    const long kbytes = nb - kb0 < slab_kbytes ? nb - kb0 : slab_kbytes;
    const long kuse = round_up(kbytes * 4, 128);
    expand_2bit_launch(d_x, rows, nb, kb0, kbytes, is_plink_format, X8, rows_pad, kpad, sms, stream);
    count_launch();
    MB_LAUNCH_CHECK();
    MB_CHECK(crossprod_tc_launch(X8, rows_pad, kuse, kpad, d_tiles, plan.num_tiles, C32, rows_pad, slab > 0,
                                 sms, stream, nullptr, 0, plan.pair));
    if (executed_ops) *executed_ops += 2.0 * 128.0 * 256.0 * (double)kuse * (double)plan.real_tiles;
  }
  return 0;  // `scratch` frees stream-ordered: the slab outlives the kernels enqueued above
}

// d_m[0..rows) = M 1, d_m[rows] = sum of them, d_m[rows + 1] = 2 sum f (1 - f) (do_scale); d_m holds
// rows + 8 doubles.  In line on `stream`.
int crossprod_row_sums(const uint8_t *d_x, long k, long rows, int is_plink_format, const double *d_f,
                       int do_scale, double *d_m, cudaStream_t stream) {
  if (do_scale && !d_f) return fail(-3, __func__, __LINE__, "allele frequencies required for scaling");
  const long nb = ceil_div(k, 4);
  if (is_plink_format) MB_CHECK(upload_lut());
  int device = 0;
  MB_CUDA(cudaGetDevice(&device));
  unsigned int *d_s = nullptr;
  CallScratch scratch(stream);
  MB_CUDA(scratch.alloc(&d_s, sizeof(unsigned int) * input_row_sums_scratch(nb)));
  cudaEvent_t join = nullptr;
  return grm_row_sums(d_x, rows, nb, k, is_plink_format, d_f, do_scale, d_s, d_m, device, stream, false, &join);
}

}  // namespace mb200
