// stager.h -- host panels that are NOT page-locked (Julia Matrix{Float64}, Fortran arrays, numpy:
// what the reference's bindings pass, src/bindings/Julia/dgemm_compressed.jl:115-134) reach the
// device through a pinned ring owned by the handle.  cudaMemcpyAsync on pageable memory is a
// staged, host-blocking copy inside the driver (one thread, no overlap with the enqueueing of the
// next kernels); here the staging memcpy runs on a small pool of host threads, slot by slot, and
// every slot's DMA is asynchronous, so the chunked copy/compute overlap of dgemm_tc.cu survives.
// Page-locked caller memory (cudaHostAlloc / cudaHostRegister, torch pin_memory) is detected with
// cudaPointerGetAttributes and copied directly.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace mb200 {

// dst[r + c*dpitch] = src[r + c*spitch], width bytes per column, split over the copy pool
// (MB200_COPY_THREADS helper threads, default min(16, cores/2) - 1; the callers take part, and
// concurrent calls share the helpers).
void parallel_copy_2d(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                      size_t height);

bool host_is_pinned(const void *p);  // page-locked and usable for async DMA on the current device

struct Stager {
  static constexpr int kSlots = 6;
  size_t slot_bytes = 0;
  uint8_t *ring = nullptr;  // pinned, kSlots * slot_bytes, allocated by the device's worker thread
  cudaEvent_t done[kSlots] = {nullptr};
  bool busy[kSlots] = {false};
  struct Pending {  // device -> slot copy in flight; the slot still has to go to the caller's memory
    bool on = false;
    double *dst = nullptr;
    long ld = 0, rows = 0;
    int cols = 0;
  } pend[kSlots];
  int next = 0;
  int device = -1;   // set by stager_acquire
  bool parked = false;

  int init(size_t slot_bytes_);
  void destroy();
  // column-major panel blocks of doubles; ld in elements.  `st` = the stream the DMA is ordered on.
  int upload(double *d, long ld_d, const double *h, long ld_h, long rows, int cols, cudaStream_t st);
  int download(double *h, long ld_h, const double *d, long ld_d, long rows, int cols, cudaStream_t st);
  int drain();  // completes every pending download (blocks until their DMAs are done)

 private:
  int acquire(int *slot);
  int retire(int slot);
};

// Rings kept between calls: cudaHostAlloc + cudaFreeHost of a 24 MB ring cost 20-40 ms, more than a
// whole configs[1] GRM.  stager_acquire hands out an initialised ring for `device` (which must be the
// current device) -- a cached one when one is free, a new one otherwise -- and stager_release parks
// it for the next call (at most a few per device stay cached; they are never freed before exit).
Stager *stager_acquire(int device, size_t slot_bytes);
void stager_release(Stager *s);

// Copy a rows x cols column-major block between host and device, directly when the host side is
// page-locked, through the stager otherwise.  Downloads through the stager complete at drain().
int panel_h2d(Stager *sg, bool pinned, double *d, long ld_d, const double *h, long ld_h, long rows,
              int cols, cudaStream_t st);
int panel_d2h(Stager *sg, bool pinned, double *h, long ld_h, const double *d, long ld_d, long rows,
              int cols, cudaStream_t st);

}  // namespace mb200
