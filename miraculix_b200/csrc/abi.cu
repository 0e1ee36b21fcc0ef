// abi.cu -- the C-ABI of libmiraculix_b200.so (include/miraculix_b200.h).
// Host-side replacement of the reference's boundary and dispatch layers:
//   src/miraculix/5codesAPI.c:37-161   (the exported symbols)
//   src/miraculix/5codesChar.cc:165-209, :344-449  (setOptions5, plinkToSxI, *5api, free5)
//   src/cuda/cuda_utils.cu:44-248      (PRINT_LEVEL / CUDA_DEVICE handling, error printing)
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include "../../include/miraculix_b200.h"
#include <nvtx3/nvToolsExt.h>
#include <mutex>
#include "common.cuh"
#include "handle.h"
#include "multi.h"

namespace mb200 {

NvtxRange::NvtxRange(const char *name) { nvtxRangePushA(name); }
NvtxRange::~NvtxRange() { nvtxRangePop(); }

// One library-owned memory pool per device for stream-ordered scratch (crossproduct slabs, tile
// lists): keeps up to 4 GB cached between calls -- a GBLUP script calls the crossproduct
// repeatedly and the operand slab otherwise costs ~10 ms of allocation per call -- without touching
// the process-wide default pool's settings.
cudaError_t pool_alloc(void **p, size_t bytes, cudaStream_t stream) {
  static std::mutex mu;
  static cudaMemPool_t pools[64] = {nullptr};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaMallocAsync(p, bytes, stream);
  cudaMemPool_t pool;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!pools[dev]) {
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = dev;
      e = cudaMemPoolCreate(&pools[dev], &props);
      if (e != cudaSuccess) { pools[dev] = nullptr; return e; }
      uint64_t keep = 4ull << 30;
      if (const char *env = getenv("MB200_POOL_KEEP_GB")) keep = (uint64_t)(atof(env) * 1073741824.0);
      cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
      // pool memory is private to its device unless access is granted: the multi-device paths read
      // each other's scratch (packed input of the crossproduct) over NVLink
      int count = 0;
      cudaGetDeviceCount(&count);
      for (int peer = 0; peer < count; peer++) {
        int can = 0;
        if (peer == dev || cudaDeviceCanAccessPeer(&can, peer, dev) != cudaSuccess || !can) continue;
        cudaMemAccessDesc desc = {};
        desc.location.type = cudaMemLocationTypeDevice;
        desc.location.id = peer;
        desc.flags = cudaMemAccessFlagsProtReadWrite;
        if (cudaMemPoolSetAccess(pools[dev], &desc, 1) != cudaSuccess) cudaGetLastError();
      }
    }
    pool = pools[dev];
  }
  return cudaMallocFromPoolAsync(p, bytes, pool, stream);
}

std::atomic<unsigned long long> g_kernel_launches{0};

static thread_local Status t_status;
static thread_local CrossprodTiming t_cp_timing;
Status &status() { return t_status; }
void clear_status() { t_status.code = 0; t_status.msg[0] = 0; }

int fail(int code, const char *func, int line, const char *fmt, ...) {
  Status &s = status();
  s.code = code ? code : -1;
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(s.msg, sizeof(s.msg), fmt, ap);
  va_end(ap);
  // reference style (cuda_utils.cu:84-118): always print, caller decides what to do
  fprintf(stderr, "miraculix_b200: error in %s (line %d): %s\n", func, line, s.msg);
  return s.code;
}

int print_level() {
  static const int lvl = [] {
    const char *e = getenv("PRINT_LEVEL");
    return e ? atoi(e) : -1;  // quiet by default
  }();
  return lvl;
}

// PRINT_LEVEL (reference cuda_utils.cu:44-82): >= 0 prints a one-time banner and a summary line per
// entry point, > 0 adds per-call details (the reference's cyan debug_info lines).  Unset = quiet:
// stdout belongs to the host program (the reference defaults to level 0).
void print_banner_once() {
  static std::once_flag once;
  if (print_level() < 0) return;
  std::call_once(once, [] {
    printf("------------------------------------------------------------\n");
    printf("\tmiraculix_b200 - B200-native compressed-genotype library\n");
    printf("Compiled on %s %s for sm_100a\n", __DATE__, __TIME__);
    printf("------------------------------------------------------------\n");
  });
}
void debug_info(const char *fmt, ...) {  // cuda_utils.cu:55-63
  if (print_level() <= 0) return;
  va_list ap;
  va_start(ap, fmt);
  printf("\033[36m\t ");
  vprintf(fmt, ap);
  printf(" \033[37m\n");
  va_end(ap);
}

Options &options() {
  static Options o = [] {  // thread-safe one-time init
    Options v;
    if (const char *e = getenv("MB200_DGEMM_MODE")) v.dgemm_mode = atoi(e);
    if (const char *e = getenv("MB200_DGEMM_SLICES")) v.dgemm_slices = atoi(e);
    if (const char *e = getenv("MB200_CROSSPROD_MODE")) v.crossprod_mode = atoi(e);
    if (const char *e = getenv("MB200_WIDE_CLUSTER")) v.wide_cluster = atoi(e) != 0;
    if (const char *e = getenv("MB200_SMALL_LIMIT")) v.small_limit = atof(e);
    if (const char *e = getenv("MB200_PIPE_CHUNKS")) v.pipe_chunks = atoi(e);
    if (const char *e = getenv("MB200_PIPE_MIN_BYTES")) v.pipe_min_bytes = atof(e);
    return v;
  }();
  return o;
}

int device_sm_count(int device) {
  static int cache[64] = {0};
  if (device < 0 || device >= 64) return 148;
  if (!cache[device]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0)
      n = 148;
    cache[device] = n;
  }
  return cache[device];
}

// CUDA_DEVICE env selects the ordinal (reference cuda_utils.cu:188-248); unlike the reference,
// an unset CUDA_VISIBLE_DEVICES is not an error, and without CUDA_DEVICE the calling thread's
// current device is kept (so torch.cuda.set_device / cudaSetDevice callers are respected).
int select_device() {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0) {
    fail((int)(e ? e : cudaErrorNoDevice), __func__, __LINE__,
         "no usable CUDA device (%s); this library has no CPU path",
         cudaGetErrorString(e ? e : cudaErrorNoDevice));
    return -1;
  }
  int dev = -1;
  if (const char *env = getenv("CUDA_DEVICE")) dev = atoi(env);
  if (dev < 0) {
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  }
  if (dev >= count) {
    fail(-4, __func__, __LINE__, "CUDA_DEVICE=%d but only %d device(s) visible", dev, count);
    return -1;
  }
  if (cudaSetDevice(dev) != cudaSuccess) {
    fail(-4, __func__, __LINE__, "cudaSetDevice(%d) failed", dev);
    return -1;
  }
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, dev) == cudaSuccess) {
    if (p.major != 10)
      fprintf(stderr, "miraculix_b200: warning: device %d is sm_%d%d; kernels are built for sm_100a\n",
              dev, p.major, p.minor);
    if (print_level() >= 0) {  // cuda_utils.cu:204-217
      print_banner_once();
      const char *env = getenv("CUDA_DEVICE");
      if (env) printf("Environment variable CUDA_DEVICE is set to %s. ", env);
      else printf("Environment variable CUDA_DEVICE is not set. ");
      printf("Using device %s (device no %d).\n", p.name, dev);
    }
  }
  return dev;
}

Handle *as_handle(void *p) {
  Handle *h = reinterpret_cast<Handle *>(p);
  if (h && h->magic == kMultiMagic) {
    fail(-31, __func__, __LINE__,
         "this entry point needs a single-device handle (the handle is sharded over several devices: "
         "use the host-pointer calls, or mb200_get_shard for a shard's handle)");
    return nullptr;
  }
  if (!h || h->magic != kHandleMagic) {
    fail(-5, __func__, __LINE__, "invalid compressed handle %p", p);
    return nullptr;
  }
  if (cudaSetDevice(h->device) != cudaSuccess) {
    fail(-4, __func__, __LINE__, "cudaSetDevice(%d) failed", h->device);
    return nullptr;
  }
  return h;
}

void destroy_handle(Handle *h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  cudaFree(h->d_snp_major);
  cudaFree(h->d_indiv_major);
  cudaFree(h->d_f);
  cudaFree(h->d_counts);
  cudaFree(h->d_B);
  cudaFree(h->d_C);
  cudaFree(h->d_w);
  cudaFree(h->d_Bq);
  cudaFree(h->d_cg);
  if (h->h_cg) cudaFreeHost(h->h_cg);
  if (h->h_pipe) cudaFreeHost(h->h_pipe);
  if (h->h_small) cudaFreeHost(h->h_small);
  if (h->copy_stream) { cudaStreamSynchronize(h->copy_stream); cudaStreamDestroy(h->copy_stream); }
  for (int i = 0; i < Handle::kPipeMax; i++)
    if (h->ev_pipe[i]) cudaEventDestroy(h->ev_pipe[i]);
  h->stager.destroy();
  for (int i = 0; i < 2; i++)
    if (h->ev_copy[i]) cudaEventDestroy(h->ev_copy[i]);
  for (int i = 0; i < Handle::kEvRing; i++) {
    if (h->ev_begin[i]) cudaEventDestroy(h->ev_begin[i]);
    if (h->ev_end[i]) cudaEventDestroy(h->ev_end[i]);
  }
  if (h->stream) cudaStreamDestroy(h->stream);
  h->magic = 0;
  delete h;
}

// Allocate everything a handle owns (device layout + per-call scratch); no repack yet.
int create_handle(long snps, long indiv, int max_n, Handle **out, int device) {
  *out = nullptr;
  if (snps <= 0 || indiv <= 0) return fail(-2, __func__, __LINE__, "snps=%ld indiv=%ld", snps, indiv);
  if (max_n < 1) max_n = 1;
  int dev = device;
  if (dev < 0) {
    dev = select_device();
    if (dev < 0) return status().code;
  } else {
    MB_CUDA(cudaSetDevice(dev));
  }
  Handle *h = new (std::nothrow) Handle();
  if (!h) return fail(-6, __func__, __LINE__, "out of host memory");
  h->device = dev;
  h->snps = snps;
  h->indiv = indiv;
  h->max_n = max_n;
  h->pitch_s = round_up(ceil_div(indiv, 4), 128);
  h->pitch_i = round_up(ceil_div(snps, 4), 128);
  h->sm_count = device_sm_count(dev);
  const long mx = snps > indiv ? snps : indiv;
#define H_CUDA(expr)                                                                  \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) {                                                          \
      int rc = fail((int)_e, __func__, __LINE__, "%s: %s", #expr, cudaGetErrorString(_e)); \
      destroy_handle(h);                                                              \
      return rc;                                                                      \
    }                                                                                 \
  } while (0)
  H_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  H_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < Handle::kPipeMax; i++)
    H_CUDA(cudaEventCreateWithFlags(&h->ev_pipe[i], cudaEventDisableTiming));
  H_CUDA(cudaMallocHost(&h->h_pipe, 64));
  for (int i = 0; i < 2; i++) H_CUDA(cudaEventCreate(&h->ev_copy[i]));
  for (int i = 0; i < Handle::kEvRing; i++) {
    H_CUDA(cudaEventCreate(&h->ev_begin[i]));
    H_CUDA(cudaEventCreate(&h->ev_end[i]));
  }
  H_CUDA(cudaMalloc(&h->d_snp_major, (size_t)snps * h->pitch_s));
  H_CUDA(cudaMalloc(&h->d_indiv_major, (size_t)indiv * h->pitch_i));
  H_CUDA(cudaMalloc(&h->d_f, sizeof(double) * snps));
  H_CUDA(cudaMalloc(&h->d_counts, sizeof(unsigned long long) * snps));
  H_CUDA(cudaMalloc(&h->d_B, sizeof(double) * mx * max_n));
  H_CUDA(cudaMalloc(&h->d_C, sizeof(double) * mx * max_n));
  H_CUDA(cudaMalloc(&h->d_w, sizeof(double) * 8 * (max_n + 16)));
  H_CUDA(cudaMemsetAsync(h->d_counts, 0, sizeof(unsigned long long) * snps, h->stream));
#undef H_CUDA
  *out = h;
  return 0;
}

int repack_chunk(Handle *h, const uint8_t *d_chunk, long s_begin, long s_count);

// Host bed -> device layout: stream the bed through two device staging buffers in chunks of
// whole 512-SNP tiles and repack each chunk as it lands.
int repack_from_host(Handle *h, const uint8_t *plink) {
  const long bs = ceil_div(h->indiv, 4);
  long rows_per_chunk = (256L << 20) / bs / 512 * 512;
  if (rows_per_chunk < 512) rows_per_chunk = 512;
  if (rows_per_chunk > round_up(h->snps, 512)) rows_per_chunk = round_up(h->snps, 512);
  uint8_t *stage[2] = {nullptr, nullptr};
  cudaEvent_t done[2] = {nullptr, nullptr};
  int rc = 0, b = 0;
  for (int i = 0; i < 2 && !rc; i++) {  // no early return: everything acquired here is released below
    cudaError_t e = cudaMalloc(&stage[i], (size_t)rows_per_chunk * bs);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) rc = fail((int)e, __func__, __LINE__, "staging buffers: %s", cudaGetErrorString(e));
  }
  for (long s = 0; s < h->snps && rc == 0; s += rows_per_chunk, b ^= 1) {
    const long cnt = h->snps - s < rows_per_chunk ? h->snps - s : rows_per_chunk;
    cudaEventSynchronize(done[b]);  // staging buffer free again
    cudaError_t e = cudaMemcpyAsync(stage[b], plink + s * bs, (size_t)cnt * bs,
                                    cudaMemcpyHostToDevice, h->stream);
    if (e != cudaSuccess) { rc = fail((int)e, __func__, __LINE__, "H2D bed: %s", cudaGetErrorString(e)); break; }
    rc = repack_chunk(h, stage[b], s, cnt);
    cudaEventRecord(done[b], h->stream);
  }
  cudaStreamSynchronize(h->stream);
  for (int i = 0; i < 2; i++) {
    if (stage[i]) cudaFree(stage[i]);
    if (done[i]) cudaEventDestroy(done[i]);
  }
  return rc;
}

// Allele frequencies (the caller's, or counts / (2 indiv) from the repack when f_host == nullptr:
// missing genotypes count as 0 there, MXinfo.h:143), their host copy, and the row-sum maxima of
// both orientations.  Synchronises the handle's stream.
int finish_handle(Handle *h, const double *f_host) {
  h->h_f.resize(h->snps);
  if (f_host) {
    memcpy(h->h_f.data(), f_host, sizeof(double) * h->snps);
    MB_CUDA(cudaMemcpyAsync(h->d_f, f_host, sizeof(double) * h->snps, cudaMemcpyHostToDevice, h->stream));
  } else {
    MB_CHECK(freq_from_counts(h));
    MB_CUDA(cudaMemcpyAsync(h->h_f.data(), h->d_f, sizeof(double) * h->snps, cudaMemcpyDeviceToHost,
                            h->stream));
  }
  MB_CUDA(cudaStreamSynchronize(h->stream));
  return rowsum_maxima(h);
}

}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_last_status(void) { return status().code; }
const char *mb200_last_message(void) { return status().msg; }
unsigned long long mb200_kernel_launches(void) { return g_kernel_launches.load(); }

int mb200_set_dgemm_mode(int mode, int slices) {
  Options &o = options();
  if (mode == 0 || mode == 1) o.dgemm_mode = mode;
  if (slices == 0 || (slices >= 3 && slices <= 8)) o.dgemm_slices = slices;
  return 0;
}
int mb200_set_dgemm_wide_cluster(int on) {
  options().wide_cluster = on ? 1 : 0;
  return 0;
}
int mb200_set_pipeline(int chunks, double min_bytes) {
  Options &o = options();
  if (chunks >= 1) o.pipe_chunks = chunks > Handle::kPipeMax ? Handle::kPipeMax : chunks;
  if (min_bytes >= 0) o.pipe_min_bytes = min_bytes;
  return 0;
}
int mb200_set_dgemm_small_limit(double products) {
  if (products >= 0) options().small_limit = products;
  return 0;
}
int mb200_set_crossprod_mode(int mode) {
  if (mode >= 0 && mode <= 2) options().crossprod_mode = mode;
  return 0;
}

void setOptions_compressed(int use_gpu, int cores, int floatLoop, int meanSubstract,
                           int ignore_missings, int do_not_center, int do_normalize,
                           int use_miraculix_freq, int variant, int print_details) {
  (void)floatLoop; (void)meanSubstract;
  Options &o = options();
  // same admissible combination as the reference GPU path (5codesChar.cc:192-193): external
  // frequencies, missings decoded as 0, no normalising; violation = message + exit like ERR0.
  if (use_miraculix_freq || !ignore_missings || do_normalize) {
    fprintf(stderr,
            "miraculix_b200: in case of 'gpu' the Fortran frequency must always be used; "
            "missings/centering/normalizing cannot treated.\n");
    exit(EXIT_FAILURE);
  }
  if (cores == 0) {
    const char *s = getenv("OMP_NUM_THREADS");
    if (s) cores = atoi(s);
    if (cores <= 0) cores = 4;  // 5codesChar.cc:177-187
  }
  o.use_gpu = use_gpu;
  o.cores = cores;
  o.centered = !do_not_center;
  o.variant = variant < 32 ? 32 : variant > 256 ? 256 : variant;
  o.print_details = print_details;
  if (print_details > 0 || print_level() > 0) {
    printf("miraculix_b200 options: gpu=always centered=%d cores=%d (host threads unused on the "
           "GPU path)%s\n", o.centered, o.cores,
           use_gpu ? "" : "  [use_gpu=0 requested: this build has no CPU path]");
  }
}

int mb200_plink2compressed_device(const unsigned char *d_plink, long snps, long indiv,
                                  const double *d_f, int max_n, void **compressed) {
  clear_status();
  if (!compressed) return fail(-2, __func__, __LINE__, "compressed == NULL");
  *compressed = nullptr;
  {
    std::vector<int> devs;
    MB_CHECK(multi_device_list(&devs));
    if (devs.size() > 1) {  // sharded: every device copies its SNP rows of the bed from where it lives
      MultiInput in;
      in.src = MultiSource::DeviceBed;
      in.plink = d_plink;
      in.f = d_f;
      return multi_create(devs, snps, indiv, max_n, in, compressed);
    }
  }
  Handle *h = nullptr;
  MB_CHECK(create_handle(snps, indiv, max_n, &h));
  int rc = repack_from_device(h, d_plink);
  if (!rc && d_f) {
    cudaError_t e = cudaMemcpyAsync(h->d_f, d_f, sizeof(double) * snps, cudaMemcpyDeviceToDevice, h->stream);
    if (e == cudaSuccess) {
      h->h_f.resize(snps);
      e = cudaMemcpyAsync(h->h_f.data(), h->d_f, sizeof(double) * snps, cudaMemcpyDeviceToHost, h->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) rc = fail((int)e, __func__, __LINE__, "%s", cudaGetErrorString(e));
    if (!rc) rc = rowsum_maxima(h);
  } else if (!rc) {
    // d_f == NULL: f = allele counts / (2 indiv) with missing genotypes counted as 0 (the repack's
    // counts; read_bed's popcount rule differs on missings and is what mb200_read_bed applies)
    rc = finish_handle(h, nullptr);
  }
  if (rc) { destroy_handle(h); return rc; }
  *compressed = h;
  return 0;
}

void plink2compressed(char *plink, char *plink_transposed, int snps, int indiv, double *f,
                      int max_n, void **compressed) {
  MB_RANGE();
  (void)plink_transposed;  // both orientations are produced on the GPU from `plink`
  clear_status();
  if (!compressed) { fail(-2, __func__, __LINE__, "compressed == NULL"); return; }
  *compressed = nullptr;
  if (!plink || !f) { fail(-2, __func__, __LINE__, "plink or f is NULL"); return; }
  std::vector<int> devs;
  if (multi_device_list(&devs)) return;
  if (devs.size() > 1) {  // MIRACULIX_NGPUS / CUDA_DEVICE list: SNP blocks sharded over the devices
    MultiInput in;
    in.src = MultiSource::HostBed;
    in.plink = reinterpret_cast<const uint8_t *>(plink);
    in.f = f;
    multi_create(devs, snps, indiv, max_n, in, compressed);
    return;
  }
  Handle *h = nullptr;
  if (create_handle(snps, indiv, max_n, &h, devs.empty() ? -1 : devs[0])) return;
  int rc = repack_from_host(h, reinterpret_cast<const uint8_t *>(plink));
  if (!rc) rc = finish_handle(h, f);
  if (rc) { destroy_handle(h); return; }
  if (print_level() >= 0)
    printf("miraculix_b200: %d SNPs x %d individuals resident on device %d (%.1f MB)\n", snps,
           indiv, h->device, ((double)snps * h->pitch_s + (double)indiv * h->pitch_i) / 1e6);
  *compressed = h;
}

// Handle from the on-device simulator (benchmarks, tests: inputs of configs[2..4] size never cross
// PCIe); one device or sharded like plink2compressed.  The matrix depends on (seed, snps, indiv)
// only, not on the number of devices.
int mb200_simulate_compressed(long snps, long indiv, unsigned long long seed, int max_n,
                              void **compressed) {
  MB_RANGE();
  clear_status();
  if (!compressed) return fail(-2, __func__, __LINE__, "compressed == NULL");
  *compressed = nullptr;
  std::vector<int> devs;
  MB_CHECK(multi_device_list(&devs));
  if (devs.size() > 1) {
    MultiInput in;
    in.src = MultiSource::Simulated;
    in.seed = seed;
    return multi_create(devs, snps, indiv, max_n, in, compressed);
  }
  Handle *h = nullptr;
  MB_CHECK(create_handle(snps, indiv, max_n, &h, devs.empty() ? -1 : devs[0]));
  uint8_t *d_bed = nullptr;
  cudaError_t e = cudaMalloc(&d_bed, (size_t)snps * ceil_div(indiv, 4));
  int rc = e == cudaSuccess ? 0 : fail((int)e, __func__, __LINE__, "bed buffer: %s", cudaGetErrorString(e));
  if (!rc) rc = simulate_plink(d_bed, snps, indiv, seed, h->stream, 0);
  if (!rc) rc = repack_from_device(h, d_bed);
  if (!rc) rc = finish_handle(h, nullptr);
  cudaFree(d_bed);
  if (rc) { destroy_handle(h); return rc; }
  *compressed = h;
  return 0;
}

int mb200_set_devices(int n, const int *devices) {
  clear_status();
  return multi_set_devices(n, devices);
}
int mb200_shard_count(void *compressed) {
  clear_status();
  if (MultiHandle *m = as_multi(compressed)) return multi_shards(m);
  return as_handle(compressed) ? 1 : -1;
}
int mb200_get_shard(void *compressed, int g, void **shard, long *snp_begin, long *snp_count) {
  clear_status();
  if (!shard) return fail(-2, __func__, __LINE__, "shard == NULL");
  *shard = nullptr;
  if (MultiHandle *m = as_multi(compressed)) {
    Handle *h = multi_shard(m, g, snp_begin, snp_count);
    if (!h) return fail(-2, __func__, __LINE__, "shard %d of %d", g, multi_shards(m));
    *shard = h;
    return 0;
  }
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  if (g != 0) return fail(-2, __func__, __LINE__, "shard %d of 1", g);
  if (snp_begin) *snp_begin = 0;
  if (snp_count) *snp_count = h->snps;
  *shard = h;
  return 0;
}

int mb200_dgemm_compressed_device(int trans, void *compressed, int n, const double *dB, long ldb,
                                  double *dC, long ldc) {
  clear_status();
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  const long m = trans ? h->snps : h->indiv, k = trans ? h->indiv : h->snps;
  if (n < 0 || n > h->max_n || ldb < k || ldc < m)
    return fail(-2, __func__, __LINE__, "bad arguments n=%d (max_n=%d) ldb=%ld (k=%ld) ldc=%ld (m=%ld)",
                n, h->max_n, ldb, k, ldc, m);
  return dgemm_device(h, trans ? 1 : 0, n, dB, ldb, dC, ldc);
}

void *mb200_stream(void *compressed) {
  Handle *h = as_handle(compressed);
  return h ? (void *)h->stream : nullptr;
}
int mb200_dgemm_last_kernel_ms(void *compressed, float *ms) {
  Handle *h = as_handle(compressed);
  if (!h || !ms) return status().code ? status().code : -2;
  if (h->ev_count == 0) return fail(-2, __func__, __LINE__, "no dgemm_compressed call yet");
  MB_CUDA(cudaStreamSynchronize(h->stream));
  const int i = (int)((h->ev_count - 1) % Handle::kEvRing);
  MB_CUDA(cudaEventElapsedTime(ms, h->ev_begin[i], h->ev_end[i]));
  return 0;
}
int mb200_dgemm_kernel_times(void *compressed, int last, float *ms, int *trans) {
  Handle *h = as_handle(compressed);
  if (!h || !ms) return status().code ? status().code : -2;
  MB_CUDA(cudaStreamSynchronize(h->stream));
  if (last > Handle::kEvRing) last = Handle::kEvRing;
  if ((unsigned long long)last > h->ev_count) last = (int)h->ev_count;
  for (int j = 0; j < last; j++) {
    const int i = (int)((h->ev_count - last + j) % Handle::kEvRing);
    MB_CUDA(cudaEventElapsedTime(&ms[j], h->ev_begin[i], h->ev_end[i]));
    if (trans) trans[j] = h->ev_trans[i];
  }
  return last;
}
int mb200_debug_read_trace(long long *out, int n) {
  clear_status();
  cudaDeviceSynchronize();
  return dgemm_tc_read_trace(out, n);
}
int mb200_dgemm_last_precision(void *compressed, int *slices) {
  Handle *h = as_handle(compressed);
  if (!h) return -1;
  return dgemm_tc_last_precision(h, slices);
}
int mb200_sync(void *compressed) {
  if (MultiHandle *m = as_multi(compressed)) return multi_sync(m);
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  MB_CUDA(cudaStreamSynchronize(h->stream));
  return 0;
}

static int is_trans(const char *trans) {  // 5codesAPI.c:73-77
  if (*trans == 'T' || *trans == 't' || *trans == 'Y' || *trans == 'y') return 1;
  if (*trans != 'N' && *trans != 'n') exit(99);
  return 0;
}

static int dgemm_host(Handle *h, int trans, int n, const double *B, long ldb, double *C, long ldc) {
  const long m = trans ? h->snps : h->indiv, k = trans ? h->indiv : h->snps;
  if (n > h->max_n)
    return fail(-2, __func__, __LINE__, "n=%d exceeds max_n=%d given to plink2compressed", n, h->max_n);
  if (ldb < k || ldc < m)
    return fail(-2, __func__, __LINE__, "leading dimension too small: ldb=%ld (need %ld) ldc=%ld (need %ld)",
                ldb, k, ldc, m);
  if (n <= 0) return 0;
  if (dgemm_small_applies(h, n))  // latency bound: one fused kernel, one synchronisation
    return dgemm_small_host(h, trans, n, B, ldb, C, ldc);
  if (options().dgemm_mode == 1)  // tensor-core path: PCIe copies pipelined against the kernels
    return dgemm_tc_host_pipelined(h, trans, n, B, ldb, C, ldc, 0, 0);
  const bool b_pin = host_is_pinned(B), c_pin = host_is_pinned(C);
  if (!b_pin || !c_pin) MB_CHECK(h->stager.init(4u << 20));
  MB_CHECK(panel_h2d(&h->stager, b_pin, h->d_B, k, B, ldb, k, n, h->stream));
  MB_CHECK(dgemm_device(h, trans, n, h->d_B, k, h->d_C, m));
  MB_CHECK(panel_d2h(&h->stager, c_pin, C, ldc, h->d_C, m, m, n, h->stream));
  if (!c_pin) MB_CHECK(h->stager.drain());
  MB_CUDA(cudaStreamSynchronize(h->stream));
  return 0;
}

// B and/or C may live on the device (SNP-sharded callers reduce the 'N' partial on the device
// before it goes back to the host); host panels are copied, pipelined where it pays.
int mb200_dgemm_compressed_mixed(int trans, void *compressed, int n, const double *B, long ldb,
                                 int b_on_device, double *C, long ldc, int c_on_device) {
  clear_status();
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  trans = trans ? 1 : 0;
  const long m = trans ? h->snps : h->indiv, k = trans ? h->indiv : h->snps;
  if (n < 0 || n > h->max_n || ldb < k || ldc < m)
    return fail(-2, __func__, __LINE__, "bad arguments n=%d (max_n=%d) ldb=%ld (k=%ld) ldc=%ld (m=%ld)",
                n, h->max_n, ldb, k, ldc, m);
  if (n == 0) return 0;
  if (b_on_device && c_on_device) return dgemm_device(h, trans, n, B, ldb, C, ldc);
  if (options().dgemm_mode == 1)
    return dgemm_tc_host_pipelined(h, trans, n, B, ldb, C, ldc, b_on_device, c_on_device);
  // fp64 SIMT family: plain staging
  const double *dB = B;
  double *dC = C;
  const bool b_pin = b_on_device || host_is_pinned(B), c_pin = c_on_device || host_is_pinned(C);
  if (!b_pin || !c_pin) MB_CHECK(h->stager.init(4u << 20));
  if (!b_on_device) {
    MB_CHECK(panel_h2d(&h->stager, b_pin, h->d_B, k, B, ldb, k, n, h->stream));
    dB = h->d_B;
    ldb = k;
  }
  if (!c_on_device) dC = h->d_C;
  MB_CHECK(dgemm_device(h, trans, n, dB, ldb, dC, c_on_device ? ldc : m));
  if (!c_on_device) {
    MB_CHECK(panel_d2h(&h->stager, c_pin, C, ldc, h->d_C, m, m, n, h->stream));
    if (!c_pin) MB_CHECK(h->stager.drain());
  }
  MB_CUDA(cudaStreamSynchronize(h->stream));
  return 0;
}

void dgemm_compressed(char *trans, void *compressed, int n, double *B, int Ldb, double *C, int Ldc) {
  MB_RANGE();
  clear_status();
  const int t = is_trans(trans);
  if (MultiHandle *m = as_multi(compressed)) {
    multi_dgemm_host(m, t, n, B, Ldb, C, Ldc);
    return;
  }
  Handle *h = as_handle(compressed);
  if (!h) return;
  dgemm_host(h, t, n, B, Ldb, C, Ldc);
  debug_info("dgemm_compressed '%c' n=%d on device %d: %d digits", t ? 'T' : 'N', n, h->device, h->last_slices);
}

void free_compressed(void **compressed) {
  MB_RANGE();
  clear_status();
  if (!compressed || !*compressed) return;
  if (MultiHandle *m = as_multi(*compressed)) {
    multi_destroy(m);
    *compressed = nullptr;
    return;
  }
  Handle *h = reinterpret_cast<Handle *>(*compressed);
  if (h->magic != kHandleMagic) { fail(-5, __func__, __LINE__, "invalid handle"); return; }
  destroy_handle(h);
  *compressed = nullptr;
}

void get_compressed_freq(void *compressed, double *f) {
  clear_status();
  if (MultiHandle *m = as_multi(compressed)) {
    if (f) multi_get_freq(m, f);
    return;
  }
  Handle *h = as_handle(compressed);
  if (!h || !f) return;
  memcpy(f, h->h_f.data(), sizeof(double) * h->snps);
}

int mb200_get_layout(void *compressed, mb200_layout *out) {
  clear_status();
  Handle *h = as_handle(compressed);
  if (!h || !out) return status().code ? status().code : -2;
  out->snps = h->snps;
  out->indiv = h->indiv;
  out->pitch_snp_major = h->pitch_s;
  out->pitch_indiv_major = h->pitch_i;
  out->d_snp_major = h->d_snp_major;
  out->d_indiv_major = h->d_indiv_major;
  out->d_f = h->d_f;
  out->d_allele_counts = h->d_counts;
  out->device = h->device;
  return 0;
}

int mb200_download_layout(void *compressed, unsigned char *snp_major, unsigned char *indiv_major,
                          unsigned long long *allele_counts) {
  clear_status();
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  MB_CUDA(cudaStreamSynchronize(h->stream));
  if (snp_major)
    MB_CUDA(cudaMemcpy(snp_major, h->d_snp_major, (size_t)h->snps * h->pitch_s, cudaMemcpyDeviceToHost));
  if (indiv_major)
    MB_CUDA(cudaMemcpy(indiv_major, h->d_indiv_major, (size_t)h->indiv * h->pitch_i, cudaMemcpyDeviceToHost));
  if (allele_counts)
    MB_CUDA(cudaMemcpy(allele_counts, h->d_counts, sizeof(unsigned long long) * h->snps, cudaMemcpyDeviceToHost));
  return 0;
}

int mb200_simulate_plink_device(unsigned char *d_plink, long snps, long indiv,
                                unsigned long long seed) {
  clear_status();
  MB_CHECK(simulate_plink(d_plink, snps, indiv, seed, 0));
  MB_CUDA(cudaDeviceSynchronize());
  return 0;
}

int mb200_simulate_plink_rows_device(unsigned char *d_plink, long snps, long indiv,
                                     unsigned long long seed, long snp_offset) {
  clear_status();
  MB_CHECK(simulate_plink(d_plink, snps, indiv, seed, 0, snp_offset));
  MB_CUDA(cudaDeviceSynchronize());
  return 0;
}

int mb200_crossprod_device(const unsigned char *d_x, long k, long rows, int is_plink_format,
                           int mode, const double *d_f, int do_scale, double *d_ans) {
  clear_status();
  return crossprod_device(d_x, k, rows, is_plink_format, mode, d_f, do_scale, d_ans, 0,
                          &t_cp_timing);
}

long mb200_crossprod_partition_rows(long rows, int part, int nparts) {
  return crossprod_partition_rows(rows, part, nparts);
}
int mb200_crossprod_partition_device(const unsigned char *d_x, long k, long rows, int is_plink_format,
                                     int mode, const double *d_f, int do_scale, int part, int nparts,
                                     double *d_panel) {
  clear_status();
  return crossprod_partition_device(d_x, k, rows, is_plink_format, mode, d_f, do_scale, part, nparts,
                                    d_panel, 0, &t_cp_timing);
}
// host-pointer convenience wrapper of the above (tests, small problems)
int mb200_grm_partition(unsigned char *snp_matrix, int snps, int indiv, bool is_plink_format, int mode,
                        const double *f, bool do_scale, int part, int nparts, double *panel) {
  clear_status();
  if (!snp_matrix || !panel || snps <= 0 || indiv <= 0)
    return fail(-2, __func__, __LINE__, "bad arguments") ? 1 : 1;
  if (select_device() < 0) return 1;
  const long nb = ceil_div(snps, 4), lrows = crossprod_partition_rows(indiv, part, nparts);
  if (lrows == 0) return 0;
  uint8_t *d_x = nullptr;
  double *d_p = nullptr, *d_f = nullptr;
  int rc = 0;
  cudaError_t e;
#define P_CUDA(expr)                                                                     \
  if (!rc && (e = (expr)) != cudaSuccess)                                                 \
    rc = fail((int)e, __func__, __LINE__, "%s: %s", #expr, cudaGetErrorString(e))
  P_CUDA(cudaMalloc(&d_x, (size_t)indiv * nb));
  P_CUDA(cudaMalloc(&d_p, sizeof(double) * lrows * indiv));
  P_CUDA(cudaMemcpy(d_x, snp_matrix, (size_t)indiv * nb, cudaMemcpyHostToDevice));
  if (f && mode) {
    P_CUDA(cudaMalloc(&d_f, sizeof(double) * snps));
    P_CUDA(cudaMemcpy(d_f, f, sizeof(double) * snps, cudaMemcpyHostToDevice));
  }
  if (!rc) rc = crossprod_partition_device(d_x, snps, indiv, is_plink_format, mode, d_f, do_scale, part,
                                           nparts, d_p, 0, &t_cp_timing);
  P_CUDA(cudaMemcpy(panel, d_p, sizeof(double) * lrows * indiv, cudaMemcpyDeviceToHost));
#undef P_CUDA
  cudaFree(d_x);
  cudaFree(d_p);
  cudaFree(d_f);
  return rc ? 1 : 0;
}

int mb200_crossprod_last_timing(double out[3]) {
  out[0] = t_cp_timing.gemm_ms;
  out[1] = t_cp_timing.epilogue_ms;
  out[2] = t_cp_timing.executed_ops;
  return 0;
}

// Host-pointer crossproduct.  Counts and GRM: the tiled out-of-core path over the selected devices
// (crossprod_host.cu; device memory bounded by TILE_SIZE / MB200_GRM_ACC_GB, not by rows^2).  LD
// needs the finished diagonal before it can scale, so it stays in core on one device.
static int crossprod_host(const unsigned char *x, int k, int rows, double *ans, int is_plink,
                          int mode, const double *f, int do_scale) {
  MB_RANGE();
  clear_status();
  if (!x || !ans || k <= 0 || rows <= 0)
    return fail(-2, __func__, __LINE__, "bad arguments (k=%d rows=%d)", k, rows) ? 1 : 1;
  std::vector<int> devs;
  if (multi_device_list(&devs)) return 1;
  if (mode != 2 && options().crossprod_mode >= 1) {  // (the SIMT cross-check family stays in core)
    const int rc = crossprod_host_tiled(x, k, rows, ans, is_plink, mode, f, do_scale, devs, &t_cp_timing);
    if (!rc && print_level() >= 0)
      printf("miraculix_b200: crossproduct %d x %d on %d device(s): device pipeline %.2f ms\n", rows, k,
             devs.empty() ? 1 : (int)devs.size(), t_cp_timing.gemm_ms);
    return rc ? 1 : 0;
  }
  if (devs.size() >= 1) {
    if (cudaSetDevice(devs[0]) != cudaSuccess) return fail(-4, __func__, __LINE__, "cudaSetDevice(%d) failed", devs[0]) ? 1 : 1;
  } else if (select_device() < 0) {
    return 1;
  }
  const long nb = ceil_div(k, 4);
  uint8_t *d_x = nullptr;
  double *d_ans = nullptr, *d_f = nullptr;
  cudaStream_t st = nullptr;
  int rc = 0;
  cudaError_t e;
#define X_CUDA(expr)                                                                     \
  if (!rc && (e = (expr)) != cudaSuccess)                                                 \
    rc = fail((int)e, __func__, __LINE__, "%s: %s", #expr, cudaGetErrorString(e))
  X_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  X_CUDA(pool_alloc(&d_x, (size_t)rows * nb, st));
  X_CUDA(pool_alloc(&d_ans, sizeof(double) * rows * rows, st));
  X_CUDA(cudaMemcpyAsync(d_x, x, (size_t)rows * nb, cudaMemcpyHostToDevice, st));
  if (f && mode) {
    const long nf = mode == 1 ? k : rows;
    X_CUDA(pool_alloc(&d_f, sizeof(double) * nf, st));
    X_CUDA(cudaMemcpyAsync(d_f, f, sizeof(double) * nf, cudaMemcpyHostToDevice, st));
  }
  if (!rc) rc = crossprod_device(d_x, k, rows, is_plink, mode, d_f, do_scale, d_ans, st, &t_cp_timing);
  X_CUDA(cudaMemcpyAsync(ans, d_ans, sizeof(double) * rows * rows, cudaMemcpyDeviceToHost, st));
  X_CUDA(cudaStreamSynchronize(st));
#undef X_CUDA
  if (d_x) cudaFreeAsync(d_x, st);
  if (d_ans) cudaFreeAsync(d_ans, st);
  if (d_f) cudaFreeAsync(d_f, st);
  if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
  return rc ? 1 : 0;
}

int snp_multiply_gpu(unsigned char *snp_matrix, int snps, int indiv, double *ans,
                     bool is_plink_format) {
  return crossprod_host(snp_matrix, snps, indiv, ans, is_plink_format, 0, nullptr, 0);
}
int mb200_grm(unsigned char *snp_matrix, int snps, int indiv, double *ans, bool is_plink_format,
              const double *f, bool do_scale) {
  return crossprod_host(snp_matrix, snps, indiv, ans, is_plink_format, 1, f, do_scale);
}
int mb200_ld(unsigned char *snp_matrix, int snps, int indiv, double *ans, bool is_plink_format,
             const double *f) {
  return crossprod_host(snp_matrix, snps, indiv, ans, is_plink_format, 2, f, 0);
}

}  // extern "C"
