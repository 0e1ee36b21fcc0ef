// stager.cu -- pinned staging ring + host copy pool for pageable caller memory (see stager.h).
#include "stager.h"
#include <stdlib.h>
#include <string.h>
#if defined(__x86_64__)
#include <emmintrin.h>
#endif
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>
#include "common.cuh"

namespace mb200 {

// ---------------------------------------------------------------------------------------------
// Copy pool: helper threads that take pieces of 2-D copies.  Several device worker threads
// (multi.cu, crossprod_host.cu) stage at the same time: every call registers its copy as a job,
// works on it itself, and idle helpers join whichever registered job still has pieces left, so
// eight concurrent callers and fifteen helpers keep about twenty cores copying (a result of
// 80 GB leaving eight devices is bound by the host memcpy otherwise, not by the links).
// ---------------------------------------------------------------------------------------------
namespace {

struct CopyJob {
  uint8_t *dst = nullptr;
  const uint8_t *src = nullptr;
  size_t dpitch = 0, spitch = 0, width = 0, height = 0;
  size_t pieces = 0, piece_bytes = 0;  // the flattened width*height bytes are cut into pieces
  std::atomic<size_t> next{0};
  int helpers_in = 0;                  // helpers working on this job (guarded by the pool mutex)
  bool streaming = false;              // non-temporal stores (large copies)
};

// Large copies bypass the cache on the store side: a staged panel is written once and read next by
// the DMA engine or, much later, by the caller, and a regular store first READS the destination line
// (write-allocate) -- three DRAM transfers per byte instead of two, on copies that run at the host's
// memory bandwidth (the pageable configs[4] GRM stages 105 GB).  SSE2 streaming stores are part of
// x86-64; other hosts use memcpy.
#if defined(__x86_64__)
void stream_copy(uint8_t *d, const uint8_t *s, size_t n) {
  if (n < 512) {
    memcpy(d, s, n);
    return;
  }
  const size_t head = (size_t)(-(intptr_t)d) & 63;  // up to the next 64-byte line of the destination
  if (head) {
    memcpy(d, s, head);
    d += head;
    s += head;
    n -= head;
  }
  const size_t body = n & ~(size_t)63;
  for (size_t i = 0; i < body; i += 64) {
    const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i));
    const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i + 16));
    const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i + 32));
    const __m128i e = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i + 48));
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + i), a);
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + i + 16), b);
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + i + 32), c);
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + i + 48), e);
  }
  if (n > body) memcpy(d + body, s + body, n - body);
}
inline void stream_fence() { _mm_sfence(); }
#else
void stream_copy(uint8_t *d, const uint8_t *s, size_t n) { memcpy(d, s, n); }
inline void stream_fence() {}
#endif

void copy_piece(const CopyJob &j, size_t piece) {
  size_t b0 = piece * j.piece_bytes;
  const size_t total = j.width * j.height;
  size_t b1 = b0 + j.piece_bytes < total ? b0 + j.piece_bytes : total;
  while (b0 < b1) {  // a piece may span several columns
    const size_t col = b0 / j.width, off = b0 % j.width;
    const size_t n = j.width - off < b1 - b0 ? j.width - off : b1 - b0;
    if (j.streaming) stream_copy(j.dst + col * j.dpitch + off, j.src + col * j.spitch + off, n);
    else memcpy(j.dst + col * j.dpitch + off, j.src + col * j.spitch + off, n);
    b0 += n;
  }
  if (j.streaming) stream_fence();  // the piece is globally visible before it is reported done
}

class CopyPool {
 public:
  static CopyPool &get() {
    static CopyPool pool;
    return pool;
  }
  int helpers() const { return (int)threads_.size(); }
  // Returns when every piece of the job is copied.  Any number of callers at a time.
  void run(CopyJob &job) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      jobs_.push_back(&job);
    }
    cv_.notify_all();
    work(job);  // the caller's share; returns once every piece has been TAKEN
    std::unique_lock<std::mutex> lk(mu_);
    for (size_t i = 0; i < jobs_.size(); i++)
      if (jobs_[i] == &job) {  // no helper can join from here on
        jobs_[i] = jobs_.back();
        jobs_.pop_back();
        break;
      }
    done_cv_.wait(lk, [&] { return job.helpers_in == 0; });  // the helpers inside have finished theirs
  }

 private:
  CopyPool() {
    int n = 0;
    if (const char *e = getenv("MB200_COPY_THREADS")) n = atoi(e);
    else {
      const unsigned hw = std::thread::hardware_concurrency();
      n = (int)(hw / 2);
      if (n > 16) n = 16;
      n -= 1;  // the callers copy too
    }
    if (n < 0) n = 0;
    if (n > 64) n = 64;
    for (int i = 0; i < n; i++) threads_.emplace_back([this] { loop(); });
  }
  ~CopyPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      quit_ = true;
    }
    cv_.notify_all();
    for (auto &t : threads_) t.join();
  }
  static void work(CopyJob &job) {
    for (;;) {
      const size_t p = job.next.fetch_add(1, std::memory_order_relaxed);
      if (p >= job.pieces) break;
      copy_piece(job, p);
    }
  }
  // a registered job that still has untaken pieces; starts at a rotating position so that the
  // helpers spread over concurrent jobs (mu_ held)
  CopyJob *pick() {
    const size_t n = jobs_.size();
    for (size_t i = 0; i < n; i++) {
      CopyJob *j = jobs_[(rr_ + i) % n];
      if (j->next.load(std::memory_order_relaxed) < j->pieces) {
        rr_ = (rr_ + i + 1) % n;
        return j;
      }
    }
    return nullptr;
  }
  void loop() {
    std::unique_lock<std::mutex> lk(mu_);
    for (;;) {
      CopyJob *job = nullptr;
      cv_.wait(lk, [&] { return quit_ || (job = pick()) != nullptr; });
      if (quit_) return;
      job->helpers_in++;
      lk.unlock();
      work(*job);
      lk.lock();
      if (--job->helpers_in == 0) done_cv_.notify_all();
    }
  }
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable cv_, done_cv_;
  std::vector<CopyJob *> jobs_;
  size_t rr_ = 0;
  bool quit_ = false;
};

}  // namespace

void parallel_copy_2d(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                      size_t height) {
  if (!width || !height) return;
  const size_t total = width * height;
  CopyJob job;
  job.dst = static_cast<uint8_t *>(dst);
  job.src = static_cast<const uint8_t *>(src);
  job.dpitch = dpitch;
  job.spitch = spitch;
  job.width = width;
  job.height = height;
  static const size_t stream_min = [] {  // MB200_COPY_STREAM_MIN: bytes from which stores bypass the cache (0 = never)
    const char *e = getenv("MB200_COPY_STREAM_MIN");
    return e ? (size_t)atof(e) : (size_t)(1 << 20);
  }();
  job.streaming = stream_min > 0 && total >= stream_min;
  constexpr size_t kMinPiece = 256 << 10;  // below ~256 KB per thread the hand-off costs more
  CopyPool &pool = CopyPool::get();
  const size_t parts = (size_t)pool.helpers() + 1;
  if (parts == 1 || total < 2 * kMinPiece) {
    job.pieces = 1;
    job.piece_bytes = total;
    copy_piece(job, 0);
    return;
  }
  size_t pieces = total / kMinPiece;
  if (pieces > 4 * parts) pieces = 4 * parts;
  job.piece_bytes = ((total + pieces - 1) / pieces + 63) & ~(size_t)63;
  job.pieces = (total + job.piece_bytes - 1) / job.piece_bytes;
  pool.run(job);
}

bool host_is_pinned(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

// ---------------------------------------------------------------------------------------------
// Stager
// ---------------------------------------------------------------------------------------------
int Stager::init(size_t slot_bytes_) {
  if (ring) return 0;
  slot_bytes = slot_bytes_;
  MB_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&ring), kSlots * slot_bytes, cudaHostAllocPortable));
  for (int i = 0; i < kSlots; i++) MB_CUDA(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
  return 0;
}

void Stager::destroy() {
  if (ring) cudaFreeHost(ring);
  ring = nullptr;
  for (int i = 0; i < kSlots; i++) {
    if (done[i]) cudaEventDestroy(done[i]);
    done[i] = nullptr;
    busy[i] = false;
    pend[i].on = false;
  }
}

// The slot's last DMA is over; a pending download leaves for the caller's memory.
int Stager::retire(int slot) {
  if (!busy[slot]) return 0;
  MB_CUDA(cudaEventSynchronize(done[slot]));
  busy[slot] = false;
  Pending &p = pend[slot];
  if (p.on) {
    p.on = false;
    parallel_copy_2d(p.dst, sizeof(double) * p.ld, ring + (size_t)slot * slot_bytes,
                     sizeof(double) * p.rows, sizeof(double) * p.rows, (size_t)p.cols);
  }
  return 0;
}

int Stager::acquire(int *slot) {
  const int s = next;
  next = (next + 1) % kSlots;
  MB_CHECK(retire(s));
  *slot = s;
  return 0;
}

int Stager::drain() {
  for (int i = 0; i < kSlots; i++) MB_CHECK(retire((next + i) % kSlots));  // oldest first
  return 0;
}

int Stager::upload(double *d, long ld_d, const double *h, long ld_h, long rows, int cols,
                   cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  const long slot_rows = (long)(slot_bytes / sizeof(double));
  const long rp_max = rows < slot_rows ? rows : slot_rows;
  for (long r0 = 0; r0 < rows; r0 += rp_max) {
    const long rp = rows - r0 < rp_max ? rows - r0 : rp_max;
    int cps = (int)(slot_rows / rp);  // whole columns per slot
    if (cps < 1) cps = 1;
    for (int c0 = 0; c0 < cols; c0 += cps) {
      const int nc = cols - c0 < cps ? cols - c0 : cps;
      int slot = 0;
      MB_CHECK(acquire(&slot));
      uint8_t *buf = ring + (size_t)slot * slot_bytes;
      parallel_copy_2d(buf, sizeof(double) * rp, h + r0 + (size_t)c0 * ld_h, sizeof(double) * ld_h,
                       sizeof(double) * rp, (size_t)nc);
      MB_CUDA(cudaMemcpy2DAsync(d + r0 + (size_t)c0 * ld_d, sizeof(double) * ld_d, buf,
                                sizeof(double) * rp, sizeof(double) * rp, (size_t)nc,
                                cudaMemcpyHostToDevice, st));
      MB_CUDA(cudaEventRecord(done[slot], st));
      busy[slot] = true;
    }
  }
  return 0;
}

int Stager::download(double *h, long ld_h, const double *d, long ld_d, long rows, int cols,
                     cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  const long slot_rows = (long)(slot_bytes / sizeof(double));
  const long rp_max = rows < slot_rows ? rows : slot_rows;
  for (long r0 = 0; r0 < rows; r0 += rp_max) {
    const long rp = rows - r0 < rp_max ? rows - r0 : rp_max;
    int cps = (int)(slot_rows / rp);
    if (cps < 1) cps = 1;
    for (int c0 = 0; c0 < cols; c0 += cps) {
      const int nc = cols - c0 < cps ? cols - c0 : cps;
      int slot = 0;
      MB_CHECK(acquire(&slot));  // may complete an older download of this ring first
      uint8_t *buf = ring + (size_t)slot * slot_bytes;
      MB_CUDA(cudaMemcpy2DAsync(buf, sizeof(double) * rp, d + r0 + (size_t)c0 * ld_d,
                                sizeof(double) * ld_d, sizeof(double) * rp, (size_t)nc,
                                cudaMemcpyDeviceToHost, st));
      MB_CUDA(cudaEventRecord(done[slot], st));
      busy[slot] = true;
      pend[slot].on = true;
      pend[slot].dst = h + r0 + (size_t)c0 * ld_h;
      pend[slot].ld = ld_h;
      pend[slot].rows = rp;
      pend[slot].cols = nc;
    }
  }
  return 0;
}

namespace {
std::mutex g_park_mu;
std::vector<Stager *> g_parked;  // idle rings, any device
constexpr size_t kMaxParkedPerDevice = 2;
}  // namespace

Stager *stager_acquire(int device, size_t slot_bytes) {
  {
    std::lock_guard<std::mutex> lk(g_park_mu);
    for (size_t i = 0; i < g_parked.size(); i++) {
      Stager *s = g_parked[i];
      if (s->device == device && s->slot_bytes == slot_bytes) {
        g_parked[i] = g_parked.back();
        g_parked.pop_back();
        s->parked = false;
        return s;
      }
    }
  }
  Stager *s = new Stager();
  s->device = device;
  if (s->init(slot_bytes) != 0) {
    s->destroy();
    delete s;
    return nullptr;
  }
  return s;
}

void stager_release(Stager *s) {
  if (!s) return;
  s->drain();  // nothing of this call stays in flight
  {
    std::lock_guard<std::mutex> lk(g_park_mu);
    size_t same = 0;
    for (Stager *p : g_parked) same += p->device == s->device;
    if (same < kMaxParkedPerDevice) {
      s->parked = true;
      s->next = 0;
      g_parked.push_back(s);
      return;
    }
  }
  s->destroy();
  delete s;
}

static cudaError_t copy2d_direct(void *dst, long dld, const void *src, long sld, long rows, int cols,
                                 cudaMemcpyKind kind, cudaStream_t st) {
  if (dld == rows && sld == rows)
    return cudaMemcpyAsync(dst, src, sizeof(double) * rows * cols, kind, st);
  return cudaMemcpy2DAsync(dst, sizeof(double) * dld, src, sizeof(double) * sld, sizeof(double) * rows,
                           (size_t)cols, kind, st);
}

int panel_h2d(Stager *sg, bool pinned, double *d, long ld_d, const double *h, long ld_h, long rows,
              int cols, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  if (pinned || !sg || !sg->ring) {
    MB_CUDA(copy2d_direct(d, ld_d, h, ld_h, rows, cols, cudaMemcpyHostToDevice, st));
    return 0;
  }
  return sg->upload(d, ld_d, h, ld_h, rows, cols, st);
}

int panel_d2h(Stager *sg, bool pinned, double *h, long ld_h, const double *d, long ld_d, long rows,
              int cols, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return 0;
  if (pinned || !sg || !sg->ring) {
    MB_CUDA(copy2d_direct(h, ld_h, d, ld_d, rows, cols, cudaMemcpyDeviceToHost, st));
    return 0;
  }
  return sg->download(h, ld_h, d, ld_d, rows, cols, st);
}

}  // namespace mb200

// Developer aid for the CPU tests of the copy pool (no GPU involved).
extern "C" int mb200_debug_parallel_copy(void *dst, size_t dpitch, const void *src, size_t spitch,
                                         size_t width, size_t height) {
  if (!dst || !src || dpitch < width || spitch < width) return -2;
  mb200::parallel_copy_2d(dst, dpitch, src, spitch, width, height);
  return 0;
}
