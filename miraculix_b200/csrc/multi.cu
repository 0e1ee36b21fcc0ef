// multi.cu -- single-process multi-device handles (see multi.h).
//
// Replaces, for more than one GPU, the single-device upload and product of the reference
// (src/cuda/dgemm_compressed_cuda.cu:43-170 plink2gpu, :218-489 dgemm_compressed_gpu): device g
// holds the SNP range S_g (boundaries multiples of 512 SNPs) in both orientations.
//
//   'T'  C[S_g, :] = (Z[:, S_g] - 2 1 f_g^T)^T B      B is replicated: device g uploads rows I_g of
//        it (1/N of the individuals) and allgather_rows_kernel pulls the other slices from the
//        peers' memory; outputs are disjoint, device g downloads its rows S_g of C.
//   'N'  C = sum_g (Z[:, S_g] - 2 1 f_g^T) B[S_g, :]   device g uploads rows S_g of B, computes
//        its partial, and reduce_partials_kernel sums rows I_g of ALL partials (peer loads, fixed
//        order: deterministic) -- a reduce-scatter; device g downloads rows I_g of C.
//
// The rigorous digit check of the tensor-core product (dgemm_tc.cu) needs the REDUCED result for
// 'N': every shard reports the bound of its partial, the reduce kernel sums the bounds and takes
// the column maxima of its slice, and every worker evaluates the same test on the host.
#include "multi.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <memory>
#include <new>
#include "../../include/miraculix_b200.h"
#include "common.cuh"
#include "team.h"

namespace mb200 {

// PRINT_LEVEL > 0: wall-clock phases of worker 0 (the reference prints its own timings the same
// way, snp_multiply_cuda.cu:354-366)
struct PhaseClock {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  bool on;
  explicit PhaseClock(bool enabled) : on(enabled && print_level() > 0) {}
  void mark(const char *what) {
    if (!on) return;
    debug_info("    %-34s %.3f ms", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  }
};

int cg_solve_device(Handle *h, const double *d_b, double *d_x, double lambda, double tol,
                    int max_iter, mb200_allreduce_fn ar, void *user, int *iters, double *resid);

constexpr int kMaxDev = 16;

// ---------------------------------------------------------------------------------------------
// device list
// ---------------------------------------------------------------------------------------------
static std::mutex g_dev_mu;
static std::vector<int> g_dev_override;
static bool g_dev_override_set = false;

int multi_set_devices(int n, const int *devs) {
  std::lock_guard<std::mutex> lk(g_dev_mu);
  if (n <= 0 || !devs) {
    g_dev_override.clear();
    g_dev_override_set = false;
    return 0;
  }
  if (n > kMaxDev) return fail(-2, __func__, __LINE__, "at most %d devices", kMaxDev);
  g_dev_override.assign(devs, devs + n);
  g_dev_override_set = true;
  return 0;
}

int multi_device_list(std::vector<int> *devs) {
  devs->clear();
  {
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (g_dev_override_set) *devs = g_dev_override;
  }
  if (devs->empty()) {
    const char *cd = getenv("CUDA_DEVICE");
    if (cd && strchr(cd, ',')) {  // explicit list
      for (const char *p = cd; *p;) {
        char *end = nullptr;
        const long v = strtol(p, &end, 10);
        if (end == p) break;
        devs->push_back((int)v);
        p = *end == ',' ? end + 1 : end;
        if (*end != ',' && *end != 0) break;
      }
    } else {
      const char *ng = getenv("MIRACULIX_NGPUS");
      if (!ng) ng = getenv("MB200_NGPUS");
      const int n = ng ? atoi(ng) : 0;
      if (n > 1) {
        const int base = cd ? atoi(cd) : 0;
        for (int i = 0; i < n; i++) devs->push_back(base + i);
      }
    }
  }
  // one device named by the environment: the ordinary single-device path (select_device reads
  // CUDA_DEVICE itself); a programmatic one-entry list is kept (the caller pins that device)
  if (devs->size() == 1) {
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (!g_dev_override_set) devs->clear();
  }
  if (devs->empty()) return 0;
  if ((int)devs->size() > kMaxDev) return fail(-2, __func__, __LINE__, "at most %d devices", kMaxDev);
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0)
    return fail((int)(e ? e : cudaErrorNoDevice), __func__, __LINE__,
                "no usable CUDA device (%s); this library has no CPU path",
                cudaGetErrorString(e ? e : cudaErrorNoDevice));
  for (int d : *devs)
    if (d < 0 || d >= count)
      return fail(-4, __func__, __LINE__, "device %d requested (MIRACULIX_NGPUS / CUDA_DEVICE list) but %d visible",
                  d, count);
  return 0;
}

// ---------------------------------------------------------------------------------------------
// exchange kernels: plain loads from the peers' memory (NVLink; same-device pointers when shards
// share a GPU)
// ---------------------------------------------------------------------------------------------
struct PeerSet {
  const double *ptr[kMaxDev];
  const double *err[kMaxDev];
  int n;
};

// out[r + c rows] = sum_p part_p[r0 + r + c ld]  (p = 0..n-1 in order), amaxC[c] = max_r |out|,
// errsum[c] = sum_p err_p[c].  blockIdx.y = column; one atomicMax per block (a per-element atomic on
// ncols addresses serialised the kernel: 325 us for 500k elements in the first ncu capture).
__global__ void __launch_bounds__(256)
reduce_partials_kernel(const PeerSet ps, long ld, long r0, long rows, int ncols,
                       double *__restrict__ out, unsigned long long *__restrict__ amaxC,
                       double *__restrict__ errsum) {
  const int c = blockIdx.y;
  double mx = 0.0;
  bool bad = false;
  for (long r = (long)blockIdx.x * 256 + threadIdx.x; r < rows; r += (long)gridDim.x * 256) {
    double s = 0.0;
    for (int p = 0; p < ps.n; p++) s += ps.ptr[p][r0 + r + (size_t)c * ld];
    out[r + (size_t)c * rows] = s;
    const double a = fabs(s);
    mx = fmax(mx, a);
    bad |= !(a <= 1.7976931348623157e308);
  }
  for (int o = 16; o; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (__any_sync(0xffffffffu, bad)) mx = __longlong_as_double(0x7ff8000000000000ll);  // NaN stays visible
  __shared__ unsigned long long red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = (unsigned long long)__double_as_longlong(mx);
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long m = 0;
    for (int i = 0; i < 8; i++) m = red[i] > m ? red[i] : m;  // non-negative doubles order like their bits
    atomicMax(&amaxC[c], m);
    if (blockIdx.x == 0 && errsum) {
      double e = 0.0;
      for (int p = 0; p < ps.n; p++) e += ps.err[p] ? ps.err[p][c] : 0.0;
      errsum[c] = e;
    }
  }
}

struct GatherSet {
  const double *src[kMaxDev];
  long r0[kMaxDev], rows[kMaxDev];
  int n, self;
};
// dst[r + c ld] = src_p[r + c ld] for the rows of every peer p != self (blockIdx.y = p)
__global__ void __launch_bounds__(256)
allgather_rows_kernel(const GatherSet gs, long ld, int ncols, double *__restrict__ dst) {
  const int p = blockIdx.y;
  if (p == gs.self || gs.src[p] == dst) return;
  const long rows = gs.rows[p], r0 = gs.r0[p], total = rows * ncols;
  const double *src = gs.src[p];
  for (long idx = (long)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (long)gridDim.x * 256) {
    const long o = r0 + idx % rows + (size_t)(idx / rows) * ld;
    dst[o] = src[o];
  }
}

// out[i] = sum_p buf_p[i] in fixed order (CG: every device ends with bit-identical vectors)
__global__ void __launch_bounds__(256)
allreduce_sum_kernel(const PeerSet ps, long count, double *__restrict__ out) {
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < count; i += (long)gridDim.x * 256) {
    double s = 0.0;
    for (int p = 0; p < ps.n; p++) s += ps.ptr[p][i];
    out[i] = s;
  }
}

// ---------------------------------------------------------------------------------------------
// the handle
// ---------------------------------------------------------------------------------------------
struct MultiHandle {
  uint32_t magic = kMultiMagic;
  int ndev = 0;
  long snps = 0, indiv = 0;
  int max_n = 0;
  std::vector<int> devices;
  std::vector<Handle *> shard;
  std::vector<long> s_begin, s_count;  // SNP range of every shard (multiples of 512 but the tail)
  std::vector<long> i_begin, i_count;  // individuals whose slice of C / B every shard moves
  std::unique_ptr<Team> team;
  // per device: reduced slice, error bounds, check block (device + pinned host), CG exchange
  std::vector<double *> d_red, d_err, d_xbuf;
  std::vector<unsigned long long *> d_chk;   // [max_n] amaxC bits, then [max_n] errsum doubles
  std::vector<unsigned long long *> h_chk;   // pinned mirror
  std::vector<cudaEvent_t> ev[2];            // exchange events, alternating
  std::vector<unsigned> phase;               // per worker: exchanges done so far (lockstep)
  std::vector<unsigned> cg_phase;
};

MultiHandle *as_multi(void *p) {
  MultiHandle *m = reinterpret_cast<MultiHandle *>(p);
  return m && m->magic == kMultiMagic ? m : nullptr;
}
int multi_shards(MultiHandle *m) { return m->ndev; }
long multi_snps(MultiHandle *m) { return m->snps; }
long multi_indiv(MultiHandle *m) { return m->indiv; }
Handle *multi_shard(MultiHandle *m, int g, long *s_begin, long *s_count) {
  if (g < 0 || g >= m->ndev) return nullptr;
  if (s_begin) *s_begin = m->s_begin[g];
  if (s_count) *s_count = m->s_count[g];
  return m->shard[g];
}
void multi_get_freq(MultiHandle *m, double *f) {
  for (int g = 0; g < m->ndev; g++)
    memcpy(f + m->s_begin[g], m->shard[g]->h_f.data(), sizeof(double) * m->s_count[g]);
}

static int enable_peers(const MultiHandle *m, int g) {
  for (int p = 0; p < m->ndev; p++) {
    if (m->devices[p] == m->devices[g]) continue;
    int can = 0;
    MB_CUDA(cudaDeviceCanAccessPeer(&can, m->devices[g], m->devices[p]));
    if (!can)
      return fail(-30, __func__, __LINE__,
                  "device %d cannot access device %d: single-process multi-GPU handles need peer access",
                  m->devices[g], m->devices[p]);
    const cudaError_t e = cudaDeviceEnablePeerAccess(m->devices[p], 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    else if (e != cudaSuccess)
      return fail((int)e, __func__, __LINE__, "cudaDeviceEnablePeerAccess(%d): %s", m->devices[p],
                  cudaGetErrorString(e));
  }
  return 0;
}

void multi_destroy(MultiHandle *m) {
  if (!m) return;
  if (m->team) {
    m->team->run([m](int g) {
      cudaSetDevice(m->devices[g]);
      if (m->shard[g]) {
        cudaStreamSynchronize(m->shard[g]->stream);
        destroy_handle(m->shard[g]);
        m->shard[g] = nullptr;
      }
      cudaFree(m->d_red[g]);
      cudaFree(m->d_err[g]);
      cudaFree(m->d_chk[g]);
      cudaFree(m->d_xbuf[g]);
      if (m->h_chk[g]) cudaFreeHost(m->h_chk[g]);
      for (int k = 0; k < 2; k++)
        if (m->ev[k][g]) cudaEventDestroy(m->ev[k][g]);
      return 0;
    });
    m->team.reset();
  }
  m->magic = 0;
  delete m;
}

int multi_create(const std::vector<int> &devs, long snps, long indiv, int max_n, const MultiInput &in,
                 void **out) {
  const MultiSource src = in.src;
  const uint8_t *plink = in.plink;
  const double *f = in.f;
  const unsigned long long seed = in.seed;
  *out = nullptr;
  if (snps <= 0 || indiv <= 0) return fail(-2, __func__, __LINE__, "snps=%ld indiv=%ld", snps, indiv);
  if (max_n < 1) max_n = 1;
  MultiHandle *m = new (std::nothrow) MultiHandle();
  if (!m) return fail(-6, __func__, __LINE__, "out of host memory");
  // SNP blocks of 512 dealt contiguously, balanced to within one block; never more shards than blocks
  const long nblocks = ceil_div(snps, 512);
  m->ndev = (int)((long)devs.size() < nblocks ? (long)devs.size() : nblocks);
  m->devices.assign(devs.begin(), devs.begin() + m->ndev);
  m->snps = snps;
  m->indiv = indiv;
  m->max_n = max_n;
  const int N = m->ndev;
  m->s_begin.resize(N);
  m->s_count.resize(N);
  m->i_begin.resize(N);
  m->i_count.resize(N);
  long b0 = 0;
  for (int g = 0; g < N; g++) {
    const long blocks = nblocks / N + (g < nblocks % N ? 1 : 0);
    const long e = (b0 + blocks) * 512 < snps ? (b0 + blocks) * 512 : snps;
    m->s_begin[g] = b0 * 512;
    m->s_count[g] = e - b0 * 512;
    b0 += blocks;
    // individual slices: multiples of 32 rows (whole 256-byte segments of a column)
    const long i0 = indiv * g / N / 32 * 32, i1 = g + 1 == N ? indiv : indiv * (g + 1) / N / 32 * 32;
    m->i_begin[g] = i0;
    m->i_count[g] = i1 - i0;
  }
  m->shard.assign(N, nullptr);
  m->d_red.assign(N, nullptr);
  m->d_err.assign(N, nullptr);
  m->d_xbuf.assign(N, nullptr);
  m->d_chk.assign(N, nullptr);
  m->h_chk.assign(N, nullptr);
  m->ev[0].assign(N, nullptr);
  m->ev[1].assign(N, nullptr);
  m->phase.assign(N, 0);
  m->cg_phase.assign(N, 0);
  m->team.reset(new Team(N));

  const long bs = ceil_div(indiv, 4);
  const int rc = m->team->run([&](int g) -> int {
    MB_CUDA(cudaSetDevice(m->devices[g]));
    MB_CHECK(enable_peers(m, g));
    Handle *h = nullptr;
    MB_CHECK(create_handle(m->s_count[g], indiv, max_n, &h, m->devices[g]));
    m->shard[g] = h;
    if (src == MultiSource::HostBed) {
      MB_CHECK(repack_from_host(h, plink + m->s_begin[g] * bs));
      MB_CHECK(finish_handle(h, f + m->s_begin[g]));
    } else if (src == MultiSource::BedFile) {
      MB_CHECK(read_bed_rows(h, in.bed_path, m->s_begin[g]));
    } else if (src == MultiSource::DeviceBed) {
      // the shard's rows of a bed that lives on some device: copied (peer or same device) into a
      // staging buffer here, repacked, frequencies from the caller (device) or from the counts
      uint8_t *d_bed = nullptr;
      MB_CUDA(cudaMalloc(&d_bed, (size_t)m->s_count[g] * bs));
      cudaError_t e = cudaMemcpyAsync(d_bed, plink + m->s_begin[g] * bs, (size_t)m->s_count[g] * bs,
                                      cudaMemcpyDefault, h->stream);
      int r = e == cudaSuccess ? 0 : fail((int)e, __func__, __LINE__, "bed shard copy: %s", cudaGetErrorString(e));
      if (!r) r = repack_from_device(h, d_bed);
      if (!r && f) {
        h->h_f.resize(m->s_count[g]);
        e = cudaMemcpyAsync(h->d_f, f + m->s_begin[g], sizeof(double) * m->s_count[g], cudaMemcpyDefault, h->stream);
        if (e == cudaSuccess)
          e = cudaMemcpyAsync(h->h_f.data(), h->d_f, sizeof(double) * m->s_count[g], cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) r = fail((int)e, __func__, __LINE__, "%s", cudaGetErrorString(e));
        if (!r) r = rowsum_maxima(h);
      } else if (!r) {
        r = finish_handle(h, nullptr);
      }
      cudaStreamSynchronize(h->stream);
      cudaFree(d_bed);
      MB_CHECK(r);
    } else {
      uint8_t *d_bed = nullptr;
      MB_CUDA(cudaMalloc(&d_bed, (size_t)m->s_count[g] * bs));
      int r = simulate_plink(d_bed, m->s_count[g], indiv, seed, h->stream, m->s_begin[g]);
      if (!r) r = repack_from_device(h, d_bed);
      if (!r) r = finish_handle(h, nullptr);
      cudaFree(d_bed);
      MB_CHECK(r);
    }
    long imax = 0;
    for (int p = 0; p < m->ndev; p++) imax = m->i_count[p] > imax ? m->i_count[p] : imax;
    MB_CUDA(cudaMalloc(&m->d_red[g], sizeof(double) * (size_t)(imax > 0 ? imax : 1) * max_n));
    MB_CUDA(cudaMalloc(&m->d_err[g], sizeof(double) * max_n));
    MB_CUDA(cudaMalloc(&m->d_chk[g], 2 * sizeof(unsigned long long) * max_n));
    MB_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&m->h_chk[g]), 2 * sizeof(unsigned long long) * max_n,
                          cudaHostAllocPortable));
    MB_CUDA(cudaMemsetAsync(m->d_err[g], 0, sizeof(double) * max_n, h->stream));
    for (int k = 0; k < 2; k++) MB_CUDA(cudaEventCreateWithFlags(&m->ev[k][g], cudaEventDisableTiming));
    MB_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
  if (rc) {
    multi_destroy(m);
    return rc;
  }
  if (print_level() > 0) {
    printf("miraculix_b200: %ld SNPs x %ld individuals sharded over %d devices:", snps, indiv, N);
    for (int g = 0; g < N; g++) printf(" %d:[%ld,%ld)", m->devices[g], m->s_begin[g], m->s_begin[g] + m->s_count[g]);
    printf("\n");
  }
  *out = m;
  return 0;
}

// Record "my part is ready", meet the peers, make my stream wait for theirs.  Every worker calls
// this the same number of times per job (lockstep), so the alternating event sets never collide.
static int exchange_point(MultiHandle *m, int g, std::vector<unsigned> &phase, cudaStream_t st) {
  const int k = phase[g]++ & 1;
  MB_CUDA(cudaEventRecord(m->ev[k][g], st));
  MB_CHECK(m->team->barrier());
  for (int p = 0; p < m->ndev; p++)
    if (p != g) MB_CUDA(cudaStreamWaitEvent(st, m->ev[k][p], 0));
  return 0;
}

static PeerSet partial_set(const MultiHandle *m) {
  PeerSet ps;
  ps.n = m->ndev;
  for (int p = 0; p < m->ndev; p++) {
    ps.ptr[p] = m->shard[p]->d_C;
    ps.err[p] = m->d_err[p];
  }
  return ps;
}

// 'N' on worker g with S digits: partial -> reduce-scatter -> my slice of C goes home
static int dgemm_n_pass(MultiHandle *m, int g, int n, const double *B, long ldb, double *C, long ldc,
                        int S) {
  Handle *h = m->shard[g];
  const bool c_pin = host_is_pinned(C);
  PhaseClock pc(g == 0);
  if (options().dgemm_mode == 1) {
    TcExternal ext;
    ext.S = S;
    ext.d_err = m->d_err[g];
    ext.no_sync = true;
    MB_CHECK(dgemm_tc_host_pipelined(h, 0, n, B + m->s_begin[g], ldb, h->d_C, m->indiv, 0, 1, &ext));
  } else {  // fp64 SIMT family: exact to fp64 rounding, no bound to check
    MB_CHECK(h->stager.init(4u << 20));
    MB_CHECK(panel_h2d(&h->stager, host_is_pinned(B), h->d_B, m->s_count[g], B + m->s_begin[g], ldb,
                       m->s_count[g], n, h->stream));
    MB_CHECK(dgemm_device(h, 0, n, h->d_B, m->s_count[g], h->d_C, m->indiv));
  }
  pc.mark("'N' partial enqueued");
  MB_CUDA(cudaMemsetAsync(m->d_chk[g], 0, 2 * sizeof(unsigned long long) * m->max_n, h->stream));
  MB_CHECK(exchange_point(m, g, m->phase, h->stream));
  pc.mark("'N' exchange point passed");
  const long rows = m->i_count[g];
  if (rows > 0) {
    long blocks = ceil_div(rows, 256);
    const long cap = ceil_div(4L * h->sm_count, n);
    if (blocks > cap) blocks = cap;
    reduce_partials_kernel<<<dim3((unsigned)blocks, (unsigned)n), 256, 0, h->stream>>>(
        partial_set(m), m->indiv, m->i_begin[g], rows, n, m->d_red[g], m->d_chk[g],
        reinterpret_cast<double *>(m->d_chk[g] + m->max_n));
    count_launch();
    MB_LAUNCH_CHECK();
    if (!c_pin) MB_CHECK(h->stager.init(4u << 20));
    MB_CHECK(panel_d2h(&h->stager, c_pin, C + m->i_begin[g], ldc, m->d_red[g], rows, rows, n, h->stream));
  }
  MB_CUDA(cudaMemcpyAsync(m->h_chk[g], m->d_chk[g], 2 * sizeof(unsigned long long) * m->max_n,
                          cudaMemcpyDeviceToHost, h->stream));
  if (!c_pin && rows > 0) MB_CHECK(h->stager.drain());
  MB_CUDA(cudaStreamSynchronize(h->stream));
  MB_CUDA(cudaStreamSynchronize(h->copy_stream));
  pc.mark("'N' streams drained");
  // nobody overwrites a partial (next pass, next call) before every peer has read it
  const int brc = m->team->barrier();
  pc.mark("'N' final barrier");
  return brc;
}

// The test of decide_kernel (dgemm_tc.cu) on the reduced result; identical on every worker.
static bool bound_fails(const MultiHandle *m, int n) {
  const double tol = dgemm_tc_tolerance();
  for (int c = 0; c < n; c++) {
    double err = 0.0, mc = 0.0;
    bool nan = false;
    for (int g = 0; g < m->ndev; g++) {
      if (m->i_count[g] <= 0) continue;
      double a, e;
      memcpy(&a, &m->h_chk[g][c], sizeof(double));
      memcpy(&e, &m->h_chk[g][m->max_n + c], sizeof(double));
      if (a != a) nan = true;
      mc = a > mc ? a : mc;
      err = e;  // the same sum on every device
    }
    if (nan || err == 0.0) continue;  // non-finite column (NaN by design) or exact zero column
    if (!(err <= tol * mc)) return true;
  }
  return false;
}

static int dgemm_n_worker(MultiHandle *m, int g, int n, const double *B, long ldb, double *C, long ldc) {
  bool adaptive = false;
  int S = 8;
  if (options().dgemm_mode == 1) S = dgemm_tc_choose_digits(n, m->indiv, m->snps, &adaptive);
  MB_CHECK(dgemm_n_pass(m, g, n, B, ldb, C, ldc, S));
  m->shard[g]->last_adaptive = adaptive;
  if (adaptive && bound_fails(m, n)) {  // every worker reaches the same verdict
    MB_CHECK(dgemm_n_pass(m, g, n, B, ldb, C, ldc, 8));
    m->shard[g]->last_slices = 8;
  }
  return 0;
}

static int dgemm_t_worker(MultiHandle *m, int g, int n, const double *B, long ldb, double *C, long ldc) {
  Handle *h = m->shard[g];
  const bool b_pin = host_is_pinned(B);
  PhaseClock pc(g == 0);
  if (!b_pin) MB_CHECK(h->stager.init(4u << 20));
  // my 1/N of the replicated B, then the peers' slices over NVLink
  MB_CHECK(panel_h2d(&h->stager, b_pin, h->d_B + m->i_begin[g], m->indiv, B + m->i_begin[g], ldb,
                     m->i_count[g], n, h->stream));
  MB_CHECK(exchange_point(m, g, m->phase, h->stream));
  if (m->ndev > 1) {
    GatherSet gs;
    gs.n = m->ndev;
    gs.self = g;
    long rmax = 1;
    for (int p = 0; p < m->ndev; p++) {
      gs.src[p] = m->shard[p]->d_B;
      gs.r0[p] = m->i_begin[p];
      gs.rows[p] = m->i_count[p];
      rmax = m->i_count[p] > rmax ? m->i_count[p] : rmax;
    }
    long bx = ceil_div(rmax * n, 256);
    if (bx > 64) bx = 64;
    allgather_rows_kernel<<<dim3((unsigned)bx, (unsigned)m->ndev), 256, 0, h->stream>>>(gs, m->indiv, n, h->d_B);
    count_launch();
    MB_LAUNCH_CHECK();
  }
  pc.mark("'T' slice up, gather enqueued");
  if (options().dgemm_mode == 1) {
    MB_CHECK(dgemm_tc_host_pipelined(h, 1, n, h->d_B, m->indiv, C + m->s_begin[g], ldc, 1, 0, nullptr));
    pc.mark("'T' product + download done");
  } else {
    const bool c_pin = host_is_pinned(C);
    if (!c_pin) MB_CHECK(h->stager.init(4u << 20));
    MB_CHECK(dgemm_device(h, 1, n, h->d_B, m->indiv, h->d_C, m->s_count[g]));
    MB_CHECK(panel_d2h(&h->stager, c_pin, C + m->s_begin[g], ldc, h->d_C, m->s_count[g], m->s_count[g], n,
                       h->stream));
    if (!c_pin) MB_CHECK(h->stager.drain());
    MB_CUDA(cudaStreamSynchronize(h->stream));
  }
  // a peer may still be pulling my slice of B: nobody leaves (and uploads the next B) before all are done
  return m->team->barrier();
}

int multi_dgemm_host(MultiHandle *m, int trans, int n, const double *B, long ldb, double *C, long ldc) {
  const long rows = trans ? m->snps : m->indiv, k = trans ? m->indiv : m->snps;
  if (n > m->max_n)
    return fail(-2, __func__, __LINE__, "n=%d exceeds max_n=%d given to plink2compressed", n, m->max_n);
  if (ldb < k || ldc < rows)
    return fail(-2, __func__, __LINE__, "leading dimension too small: ldb=%ld (need %ld) ldc=%ld (need %ld)",
                ldb, k, ldc, rows);
  if (n <= 0) return 0;
  return m->team->run([&](int g) -> int {
    MB_CUDA(cudaSetDevice(m->devices[g]));
    return trans ? dgemm_t_worker(m, g, n, B, ldb, C, ldc) : dgemm_n_worker(m, g, n, B, ldb, C, ldc);
  });
}

int multi_sync(MultiHandle *m) {
  return m->team->run([m](int g) -> int {
    MB_CUDA(cudaSetDevice(m->devices[g]));
    MB_CUDA(cudaStreamSynchronize(m->shard[g]->stream));
    return 0;
  });
}

// ---------------------------------------------------------------------------------------------
// conjugate gradients over the shards (solver.cu runs the reference's loop on every device in
// lockstep; this is its all-reduce hook: one exchange of indiv doubles per matrix-vector product)
// ---------------------------------------------------------------------------------------------
struct CgHookCtx {
  MultiHandle *m;
  int g;
};
static int cg_allreduce_hook(double *d_buf, long count, void *stream, void *user) {
  CgHookCtx *ctx = static_cast<CgHookCtx *>(user);
  MultiHandle *m = ctx->m;
  const int g = ctx->g;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the exchange buffer alternates: a peer may still read last product's copy while this one lands
  const int k = m->cg_phase[g] & 1;
  double *mine = m->d_xbuf[g] + (size_t)k * m->indiv;
  MB_CUDA(cudaMemcpyAsync(mine, d_buf, sizeof(double) * count, cudaMemcpyDeviceToDevice, st));
  MB_CHECK(exchange_point(m, g, m->cg_phase, st));
  PeerSet ps;
  ps.n = m->ndev;
  for (int p = 0; p < m->ndev; p++) {
    ps.ptr[p] = m->d_xbuf[p] + (size_t)k * m->indiv;
    ps.err[p] = nullptr;
  }
  long blocks = ceil_div(count, 256);
  if (blocks > 2L * m->shard[g]->sm_count) blocks = 2L * m->shard[g]->sm_count;
  allreduce_sum_kernel<<<(unsigned)blocks, 256, 0, st>>>(ps, count, d_buf);
  count_launch();
  MB_LAUNCH_CHECK();
  return 0;
}

int multi_cg_solve(MultiHandle *m, const double *b, double *x, double lambda, double tol,
                   int max_iter, int *iters, double *resid) {
  std::vector<int> its(m->ndev, 0);
  std::vector<double> res(m->ndev, 0.0);
  const size_t bytes = sizeof(double) * (size_t)m->indiv;
  const int rc = m->team->run([&](int g) -> int {
    MB_CUDA(cudaSetDevice(m->devices[g]));
    Handle *h = m->shard[g];
    if (!m->d_xbuf[g]) MB_CUDA(cudaMalloc(&m->d_xbuf[g], 2 * bytes));
    m->cg_phase[g] = 0;  // same on every worker: the event sets of the exchanges stay in step
    double *d = nullptr;
    MB_CUDA(cudaMalloc(&d, 2 * bytes));
    int r = 0;
    cudaError_t e = cudaMemcpyAsync(d, b, bytes, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + m->indiv, x, bytes, cudaMemcpyHostToDevice, h->stream);
    if (e != cudaSuccess) r = fail((int)e, __func__, __LINE__, "H2D: %s", cudaGetErrorString(e));
    // every worker holds a valid xbuf before anybody's first exchange reads it
    if (!r) r = m->team->barrier();
    CgHookCtx ctx{m, g};
    if (!r) r = cg_solve_device(h, d, d + m->indiv, lambda, tol, max_iter, cg_allreduce_hook, &ctx, &its[g], &res[g]);
    if (!r && g == 0) {
      e = cudaMemcpyAsync(x, d + m->indiv, bytes, cudaMemcpyDeviceToHost, h->stream);
      if (e != cudaSuccess) r = fail((int)e, __func__, __LINE__, "D2H: %s", cudaGetErrorString(e));
    }
    if (!r && cudaStreamSynchronize(h->stream) != cudaSuccess) r = fail(-1, __func__, __LINE__, "stream sync failed");
    if (!r) r = m->team->barrier();  // peers may still read my exchange buffers
    cudaFree(d);
    return r;
  });
  if (rc) return rc;
  if (iters) *iters = its[0];
  if (resid) *resid = res[0];
  return 0;
}

}  // namespace mb200
