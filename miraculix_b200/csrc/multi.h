// multi.h -- single-process multi-device handles behind the reference C ABI.
//
// The reference's callers are single-process (Julia ccall, src/bindings/Julia/dgemm_compressed.jl:
// 81-134; Fortran bind(C), src/bindings/Fortran/mod5codesapi.f90:46-69) and choose a device only
// through CUDA_DEVICE (src/cuda/cuda_utils.cu:188-248).  Here the same `void *compressed` can stand
// for a matrix sharded by 512-SNP blocks over several devices of ONE process:
//   MIRACULIX_NGPUS=N            devices CUDA_DEVICE .. CUDA_DEVICE + N - 1 (CUDA_DEVICE default 0)
//   CUDA_DEVICE=0,2,5            an explicit list
//   mb200_set_devices(n, list)   the same, programmatically
// One host thread per device (the caller's thread drives the first one), events for every
// cross-device ordering, and hand-written kernels that read the peers' memory over NVLink for the
// exchanges: 'T' needs an all-gather of the replicated B (every device uploads 1/N of it), 'N' a
// reduce-scatter of the indiv x n partials (every device downloads 1/N of C).  No panel crosses
// PCIe more than once; no NCCL dependency.
#pragma once
#include <stdint.h>
#include <vector>
#include "handle.h"

namespace mb200 {

constexpr uint32_t kMultiMagic = 0x4d42324du;  // "MB2M"
struct MultiHandle;

// Devices the next handle / crossproduct call should use: empty = one device (select_device()).
// A list may name a device more than once (the shards then share it: how the multi-device logic is
// exercised on a one-GPU box).  Returns 0, or a status code after fail().
int multi_device_list(std::vector<int> *devs);
int multi_set_devices(int n, const int *devs);

MultiHandle *as_multi(void *p);  // nullptr (no status) unless p is a live multi-device handle

enum class MultiSource { HostBed, Simulated, BedFile, DeviceBed };
struct MultiInput {
  MultiSource src = MultiSource::HostBed;
  const uint8_t *plink = nullptr;    // HostBed: host bed payload;  DeviceBed: device pointer (any device)
  const double *f = nullptr;         // HostBed: host frequencies;  DeviceBed: device pointer or nullptr
  unsigned long long seed = 0;       // Simulated
  const char *bed_path = nullptr;    // BedFile
};
int multi_create(const std::vector<int> &devs, long snps, long indiv, int max_n, const MultiInput &in,
                 void **out);
int read_bed_rows(Handle *h, const char *path, long row_offset);  // bedfile.cu
void multi_destroy(MultiHandle *m);
int multi_dgemm_host(MultiHandle *m, int trans, int n, const double *B, long ldb, double *C, long ldc);
int multi_cg_solve(MultiHandle *m, const double *b, double *x, double lambda, double tol,
                   int max_iter, int *iters, double *resid);
void multi_get_freq(MultiHandle *m, double *f);
int multi_sync(MultiHandle *m);
int multi_shards(MultiHandle *m);
Handle *multi_shard(MultiHandle *m, int g, long *s_begin, long *s_count);
long multi_snps(MultiHandle *m);
long multi_indiv(MultiHandle *m);

// Host-pointer crossproduct / GRM, out of core and over `devs` (one entry = one device): row-tile
// partitions accumulate int32 counts per device, a finishing kernel turns chunks of them into
// fp64 (with the GRM update) in both orientations and the chunks stream into the caller's
// column-major `ans`.  Replaces the host tile loop of src/cuda/snp_multiply_cuda.cu:190-340.
int crossprod_host_tiled(const uint8_t *x, long k, long rows, double *ans, int is_plink, int mode,
                         const double *f, int do_scale, const std::vector<int> &devs,
                         CrossprodTiming *timing);

}  // namespace mb200
