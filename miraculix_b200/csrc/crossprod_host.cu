// crossprod_host.cu -- host-pointer crossproduct / GRM, out of core and over several devices.
//
// Replaces the host tile loop of the reference (src/cuda/snp_multiply_cuda.cu:190-340: per tile
// pair H2D, CUTLASS launch, D2H of an int32 tile, host cast + scatter into ans[i,j] and ans[j,i])
// and keeps its property that device memory does not grow with indiv^2 (TILE_SIZE, :53-74).
//
// Device g of N (one host thread each, team.h):
//   1. uploads 1/N of the packed matrix and pulls the rest from the peers (NVLink), so the input
//      crosses PCIe once in total;
//   2. for every band it owns: accumulates the int32 counts of a band of consecutive row tiles
//      (against all column tiles at or right of the diagonal) over all K slabs; bands have about
//      equal AREA of the upper triangle and are dealt round-robin, so the devices stay balanced;
//   3. finish_chunk_kernel turns a chunk of the band's rows into fp64 -- with the exact GRM update
//      of crossproduct.jl:91-106 -- in BOTH orientations, and the chunk streams into the caller's
//      column-major matrix with two large 2-D copies: the rows of the panel are the columns of
//      `ans` from the chunk's first row down, the transposed panel fills the same columns' rows
//      to the right (the reference's ans[j,i] = ans[i,j] scatter).  Chunks are double-buffered:
//      the copy of one overlaps the finishing of the next and the GEMM of the next band.
// Pageable `ans` / `x` (Julia, Fortran arrays) go through the pinned ring of stager.h.
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <memory>
#include "common.cuh"
#include "multi.h"
#include "team.h"

namespace mb200 {

constexpr int FC_T = 32;  // 32 x 32 sub-tiles

// Chunk = global rows [c0, c0 + crows) (c0 a multiple of 256) of the band that starts at global
// row b0; counts C32[i - b0][j] (ld32) are valid for j >= 256 floor(i / 256).  For every i of the
// chunk and every column j >= c0:   P[i - c0][j - c0] (ld n - c0)  and  PT[j - c0][i - c0] (ld crows)
// receive the finished value; entries left of a row's diagonal tile (c0 <= j < 256 floor(i / 256))
// are taken from the mirrored count C32[j - b0][i], which lies in the same chunk.
__global__ void __launch_bounds__(256)
finish_chunk_kernel(const int32_t *__restrict__ C32, long ld32, long b0, long c0, long crows, long n,
                    int affine, int do_scale, const double *__restrict__ m,
                    const double *__restrict__ scal, double *__restrict__ P, double *__restrict__ PT) {
  __shared__ double t[FC_T][FC_T + 1];
  const long i0 = c0 + (long)blockIdx.y * FC_T;   // global row of sub-tile row 0
  const long j0 = c0 + (long)blockIdx.x * FC_T;   // global column of sub-tile column 0
  if (i0 >= n || j0 >= n) return;
  const long first = i0 / 256 * 256;              // same for the 32 rows of the sub-tile
  const bool mirrored = j0 < first;               // whole sub-tile left of the diagonal tile
  const CpAffine aff = cp_affine_setup(n, affine, do_scale, scal);
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const long w = n - c0;
  if (!mirrored) {
    for (int yy = ty; yy < FC_T; yy += 8) {
      const long i = i0 + yy, j = j0 + tx;
      double v = 0.0;
      if (i < n && j < n) {
        const int cnt = C32[(i - b0) * ld32 + j];
        v = affine ? cp_affine_m(cnt, m[i], m[j], aff) : (double)cnt;
        P[(i - c0) * w + (j - c0)] = v;
      }
      t[yy][tx] = v;
    }
    __syncthreads();
    for (int yy = ty; yy < FC_T; yy += 8) {
      const long j = j0 + yy, i = i0 + tx;
      if (i < n && j < n) PT[(j - c0) * crows + (i - c0)] = t[tx][yy];
    }
  } else {
    // value (i, j) = value (j, i): read C32 rows j (coalesced along i), write PT directly, P via smem
    for (int yy = ty; yy < FC_T; yy += 8) {
      const long j = j0 + yy, i = i0 + tx;
      double v = 0.0;
      if (i < n && j < n) {
        const int cnt = C32[(j - b0) * ld32 + i];
        v = affine ? cp_affine_m(cnt, m[j], m[i], aff) : (double)cnt;
        PT[(j - c0) * crows + (i - c0)] = v;
      }
      t[yy][tx] = v;
    }
    __syncthreads();
    for (int yy = ty; yy < FC_T; yy += 8) {
      const long i = i0 + yy, j = j0 + tx;
      if (i < n && j < n) P[(i - c0) * w + (j - c0)] = t[tx][yy];
    }
  }
}

namespace {

struct DevCtx {
  int device = 0;
  cudaStream_t s_main = nullptr, s_copy = nullptr;
  uint8_t *d_x = nullptr;
  double *d_f = nullptr, *d_m = nullptr;
  int32_t *C32 = nullptr;
  double *P[2] = {nullptr, nullptr}, *PT[2] = {nullptr, nullptr};
  cudaEvent_t fin[2] = {nullptr, nullptr}, freed[2] = {nullptr, nullptr}, up = nullptr;
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  Stager *stager = nullptr;  // leased per call from the parked rings (stager.h)
  std::vector<BandPlan> plans;
  double ops = 0.0;
  void release() {
    for (BandPlan &p : plans) crossprod_band_plan_free(&p, s_main);
    if (s_main) cudaStreamSynchronize(s_main);
    if (s_copy) cudaStreamSynchronize(s_copy);
    if (d_x) cudaFreeAsync(d_x, s_main);
    if (d_f) cudaFreeAsync(d_f, s_main);
    if (d_m) cudaFreeAsync(d_m, s_main);
    if (C32) cudaFreeAsync(C32, s_main);
    for (int b = 0; b < 2; b++) {
      if (P[b]) cudaFreeAsync(P[b], s_main);
      if (PT[b]) cudaFreeAsync(PT[b], s_main);
      if (fin[b]) cudaEventDestroy(fin[b]);
      if (freed[b]) cudaEventDestroy(freed[b]);
    }
    if (s_main) cudaStreamSynchronize(s_main);
    if (up) cudaEventDestroy(up);
    if (t0) cudaEventDestroy(t0);
    if (t1) cudaEventDestroy(t1);
    stager_release(stager);
    stager = nullptr;
    if (s_main) cudaStreamDestroy(s_main);
    if (s_copy) cudaStreamDestroy(s_copy);
  }
};

// flat byte upload, through the stager's ring when the source is pageable
int upload_bytes(DevCtx &c, bool pinned, uint8_t *d, const uint8_t *h, size_t bytes) {
  if (!bytes) return 0;
  if (pinned) {
    MB_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, c.s_main));
    return 0;
  }
  const size_t whole = bytes / 8;  // the ring moves doubles; the (< 8 byte) tail goes directly
  if (whole) MB_CHECK(c.stager->upload(reinterpret_cast<double *>(d), (long)whole,
                                      reinterpret_cast<const double *>(h), (long)whole, (long)whole, 1, c.s_main));
  if (bytes % 8) MB_CUDA(cudaMemcpyAsync(d + whole * 8, h + whole * 8, bytes % 8, cudaMemcpyHostToDevice, c.s_main));
  return 0;
}

}  // namespace

int crossprod_host_tiled(const uint8_t *x, long k, long rows, double *ans, int is_plink, int mode,
                         const double *f, int do_scale, const std::vector<int> &devs_in,
                         CrossprodTiming *timing) {
  if (!x || !ans || k <= 0 || rows <= 0) return fail(-2, __func__, __LINE__, "bad arguments (k=%ld rows=%ld)", k, rows);
  if (mode < 0 || mode > 1) return fail(-2, __func__, __LINE__, "mode %d not supported on the tiled path", mode);
  if (mode == 1 && (double)rows * (double)rows * 4.0 * (double)k >= 4.0e18)
    return fail(-2, __func__, __LINE__, "GRM of %ld x %ld exceeds the exact integer epilogue", rows, k);
  if (mode == 1 && do_scale && !f) return fail(-3, __func__, __LINE__, "allele frequencies required for scaling");
  std::vector<int> devs = devs_in;
  if (devs.empty()) {
    const int d = select_device();
    if (d < 0) return status().code;
    devs.push_back(d);
  }
  const long Mt = ceil_div(rows, 128);
  if ((long)devs.size() > Mt) devs.resize((size_t)Mt);
  const int N = (int)devs.size();
  const long nb = ceil_div(k, 4), rows_pad = round_up(rows, 256);

  // Bands of consecutive row tiles (even tile counts: CTA pairs, 256-row diagonal blocks).  Their
  // height is capped by TILE_SIZE (the reference's knob, rows here) and MB200_GRM_ACC_GB (bytes of
  // the int32 accumulator, default 16 GiB); within the cap they are cut to about equal AREA of the
  // upper triangle, at least MB200_GRM_PASSES (default 3) per device so that the download of one
  // band overlaps the GEMM of the next, and dealt round-robin to the devices.
  long tile_rows = 1L << 40;
  if (const char *e = getenv("TILE_SIZE")) {
    const long v = atol(e);
    if (v > 0) tile_rows = v < 256 ? 256 : v / 256 * 256;
  }
  double acc_gb = 16.0;
  if (const char *e = getenv("MB200_GRM_ACC_GB")) acc_gb = atof(e) > 0 ? atof(e) : acc_gb;
  const long mem_rows = (long)(acc_gb * 1073741824.0 / 4.0 / (double)rows_pad) / 256 * 256;
  if (mem_rows < tile_rows) tile_rows = mem_rows < 256 ? 256 : mem_rows;
  long passes = 3;
  if (const char *e = getenv("MB200_GRM_PASSES")) passes = atol(e) > 0 ? atol(e) : passes;
  const long Nt = ceil_div(rows, 256);
  std::vector<int> band_begin;  // in 128-row tiles; band b = [band_begin[b], band_begin[b + 1])
  {
    double area = 0.0;          // column tiles (of 256) at or right of the diagonal, per tile pair
    for (long pr = 0; pr < Nt; pr++) area += (double)(Nt - pr);
    const double target = area / (double)(N * passes);
    double acc = 0.0;
    band_begin.push_back(0);
    for (long pr = 0; pr < Nt; pr++) {  // pr = pair of row tiles (2 pr, 2 pr + 1)
      acc += (double)(Nt - pr);
      const long height = (pr + 1) * 256 - (long)band_begin.back() * 128;
      if (pr + 1 < Nt && (acc >= target || height + 256 > tile_rows)) {
        band_begin.push_back((int)(2 * (pr + 1)));
        acc = 0.0;
      }
    }
    band_begin.push_back((int)Mt);
  }
  const int nbands = (int)band_begin.size() - 1;
  long acc_rows = 0;  // rows of the tallest band, padded to whole CTA pairs
  for (int b = 0; b < nbands; b++) {
    const long h = round_up((long)(band_begin[b + 1] - band_begin[b]) * 128, 256);
    acc_rows = h > acc_rows ? h : acc_rows;
  }
  // chunk of band rows that is finished and downloaded at once (two orientations, two buffers):
  // at least 512 rows, so the transposed copy moves segments of >= 4 KB
  double chunk_mb = 256.0;
  if (const char *e = getenv("MB200_GRM_CHUNK_MB")) chunk_mb = atof(e) > 0 ? atof(e) : chunk_mb;
  long chunk_rows = (long)(chunk_mb * 1048576.0 / (8.0 * (double)rows)) / 256 * 256;
  const long min_chunk = getenv("MB200_GRM_CHUNK_MB") ? 256 : 512;  // an explicit setting may go lower (tests)
  if (chunk_rows < min_chunk) chunk_rows = min_chunk;
  if (chunk_rows > acc_rows) chunk_rows = acc_rows;

  const bool ans_pin = host_is_pinned(ans), x_pin = host_is_pinned(x);
  const auto w0 = std::chrono::steady_clock::now();
  auto since = [&](void) { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count(); };
  debug_info("crossproduct %ld x %ld: %d device(s), %d band(s), accumulator %ld rows, chunks of %ld rows, ans %s, x %s",
             rows, k, N, nbands, acc_rows, chunk_rows, ans_pin ? "pinned" : "pageable", x_pin ? "pinned" : "pageable");
  std::vector<DevCtx> ctx(N);
  Team team(N);
  // slices of the packed rows each device uploads itself
  std::vector<long> xr0(N + 1);
  for (int g = 0; g <= N; g++) xr0[g] = rows * g / N;

  const int rc = team.run([&](int g) -> int {
    DevCtx &c = ctx[g];
    c.device = devs[g];
    MB_CUDA(cudaSetDevice(c.device));
    for (int p = 0; p < N; p++) {  // peer access for the gather of the packed input, once per pair
      static std::atomic<unsigned long long> enabled[64];
      const int a = c.device, b = devs[p];
      if (a == b || a < 0 || a >= 64 || b < 0 || b >= 64 || (enabled[a].load() >> b & 1ull)) continue;
      int can = 0;
      MB_CUDA(cudaDeviceCanAccessPeer(&can, a, b));
      if (can) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
        if (e != cudaSuccess) cudaGetLastError();  // already enabled, or copies fall back to staging
      }
      enabled[a].fetch_or(1ull << b);
    }
    MB_CUDA(cudaStreamCreateWithFlags(&c.s_main, cudaStreamNonBlocking));
    MB_CUDA(cudaStreamCreateWithFlags(&c.s_copy, cudaStreamNonBlocking));
    for (int b = 0; b < 2; b++) {
      MB_CUDA(cudaEventCreateWithFlags(&c.fin[b], cudaEventDisableTiming));
      MB_CUDA(cudaEventCreateWithFlags(&c.freed[b], cudaEventDisableTiming));
    }
    MB_CUDA(cudaEventCreateWithFlags(&c.up, cudaEventDisableTiming));
    MB_CUDA(cudaEventCreate(&c.t0));
    MB_CUDA(cudaEventCreate(&c.t1));
    if (!ans_pin || !x_pin) {
      c.stager = stager_acquire(c.device, 4u << 20);
      if (!c.stager) return status().code ? status().code : fail(-1, __func__, __LINE__, "no pinned staging ring");
    }
    // scratch from the library's pool (cached between calls: a GBLUP script calls this repeatedly)
    // the packed matrix is read by the peers (the library pool grants them access, abi.cu)
    MB_CUDA(pool_alloc(&c.d_x, (size_t)rows * nb, c.s_main));
    MB_CUDA(pool_alloc(&c.C32, sizeof(int32_t) * (size_t)acc_rows * rows_pad, c.s_main));
    for (int b = 0; b < 2; b++) {
      MB_CUDA(pool_alloc(&c.P[b], sizeof(double) * (size_t)chunk_rows * rows, c.s_main));
      MB_CUDA(pool_alloc(&c.PT[b], sizeof(double) * (size_t)chunk_rows * rows, c.s_main));  // both: crows x (n - c0) <= this
    }
    if (mode == 1) {
      MB_CUDA(pool_alloc(&c.d_m, sizeof(double) * (rows + 8), c.s_main));
      if (f) {
        MB_CUDA(pool_alloc(&c.d_f, sizeof(double) * k, c.s_main));
        MB_CUDA(cudaMemcpyAsync(c.d_f, f, sizeof(double) * k, cudaMemcpyHostToDevice, c.s_main));
      }
    }
    // work lists of my bands, uploaded while the stream is still empty (the upload synchronises it)
    for (int b = g; b < nbands; b += N) {
      c.plans.emplace_back();
      MB_CHECK(crossprod_band_plan(rows, band_begin[b], band_begin[b + 1], c.s_main, &c.plans.back()));
    }
    if (g == 0) debug_info("  setup (streams, pool scratch, work lists) done at %.2f ms", since());
    // 1. my rows of the packed matrix from the host, the rest from the peers
    MB_CHECK(upload_bytes(c, x_pin, c.d_x + xr0[g] * nb, x + xr0[g] * nb, (size_t)(xr0[g + 1] - xr0[g]) * nb));
    MB_CUDA(cudaEventRecord(c.up, c.s_main));
    MB_CHECK(team.barrier());
    for (int p = 0; p < N; p++) {
      if (p == g || xr0[p + 1] == xr0[p]) continue;
      MB_CUDA(cudaStreamWaitEvent(c.s_main, ctx[p].up, 0));
      MB_CUDA(cudaMemcpyAsync(c.d_x + xr0[p] * nb, ctx[p].d_x + xr0[p] * nb, (size_t)(xr0[p + 1] - xr0[p]) * nb,
                              cudaMemcpyDefault, c.s_main));
    }
    if (g == 0) debug_info("  input upload + peer gather enqueued at %.2f ms", since());
    MB_CUDA(cudaEventRecord(c.t0, c.s_main));
    if (mode == 1) MB_CHECK(crossprod_row_sums(c.d_x, k, rows, is_plink, c.d_f, do_scale, c.d_m, c.s_main));
    // 2./3. my bands
    int buf = 0;
    bool used[2] = {false, false};
    size_t pi = 0;
    for (int b = g; b < nbands; b += N, pi++) {
      const int tb = band_begin[b], te = band_begin[b + 1];
      const long b0 = (long)tb * 128, b1 = (long)te * 128 < rows ? (long)te * 128 : rows;
      MB_CHECK(crossprod_band_counts(c.d_x, k, rows, is_plink, c.plans[pi], c.C32, c.s_main, &c.ops));
      if (g == 0) debug_info("  band %d [%d, %d) GEMM enqueued at %.2f ms", b, tb, te, since());
      for (long c0 = b0; c0 < b1; c0 += chunk_rows, buf ^= 1) {
        const long crows = b1 - c0 < chunk_rows ? b1 - c0 : chunk_rows;  // real rows (the last may be ragged)
        const long w = rows - c0;
        if (used[buf]) MB_CUDA(cudaStreamWaitEvent(c.s_main, c.freed[buf], 0));  // its last download is done
        dim3 grid((unsigned)ceil_div(w, FC_T), (unsigned)ceil_div(crows, FC_T));
        finish_chunk_kernel<<<grid, 256, 0, c.s_main>>>(c.C32, rows_pad, b0, c0, crows, rows, mode == 1, do_scale,
                                                        c.d_m, c.d_m ? c.d_m + rows : nullptr, c.P[buf], c.PT[buf]);
        count_launch();
        MB_LAUNCH_CHECK();
        MB_CUDA(cudaEventRecord(c.fin[buf], c.s_main));
        MB_CUDA(cudaStreamWaitEvent(c.s_copy, c.fin[buf], 0));
        // rows of the panel = columns c0.. of ans, from row c0 down (contiguous runs of n - c0)
        MB_CHECK(panel_d2h(c.stager, ans_pin, ans + c0 + c0 * rows, rows, c.P[buf], w, w, (int)crows, c.s_copy));
        // transposed panel = rows c0.. of the columns from c0 on (runs of crows)
        MB_CHECK(panel_d2h(c.stager, ans_pin, ans + c0 + c0 * rows, rows, c.PT[buf], crows, crows, (int)w, c.s_copy));
        MB_CUDA(cudaEventRecord(c.freed[buf], c.s_copy));
        used[buf] = true;
      }
    }
    MB_CUDA(cudaEventRecord(c.t1, c.s_main));
    if (g == 0) debug_info("  all work enqueued at %.2f ms", since());
    if (!ans_pin) MB_CHECK(c.stager->drain());
    MB_CUDA(cudaStreamSynchronize(c.s_main));
    MB_CUDA(cudaStreamSynchronize(c.s_copy));
    if (g == 0) debug_info("  streams drained at %.2f ms", since());
    // peers may still be pulling my slice of the input
    return team.barrier();
  });
  if (timing && !rc) {
    float ms = 0;
    cudaSetDevice(ctx[0].device);
    cudaEventElapsedTime(&ms, ctx[0].t0, ctx[0].t1);
    timing->gemm_ms = ms;
    timing->epilogue_ms = 0;
    timing->executed_ops = 0;
    for (int g = 0; g < N; g++) timing->executed_ops += ctx[g].ops;
  }
  team.run([&](int g) -> int {
    cudaSetDevice(devs[g]);
    ctx[g].release();
    return 0;
  });
  debug_info("  released at %.2f ms", since());
  return rc;
}

}  // namespace mb200
