// bedfile.cu -- PLINK .bed ingest straight into a compressed handle.
//
// Replaces the host path every reference example runs before plink2compressed:
//   read_bed                    src/bindings/Julia/read_plink.jl:161-228  (magic check, SNP / sample
//                               counts = newline counts of .bim / .fam, allele frequency = set bits
//                               per SNP / (2 indiv), :197-203)
//   transpose_genotype_matrix   src/bindings/Julia/compressed_operations.jl:45-66 (host bit loop)
// Here the file is streamed through two pinned buffers; every chunk goes H2D, is repacked into both
// device orientations (repack.cu) and has its set bits counted on the device while the next chunk
// is being read, so the bed never exists as a whole in host or device memory.
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <string>
#include "../../include/miraculix_b200.h"
#include "common.cuh"
#include "handle.h"
#include "multi.h"

namespace mb200 {

int repack_chunk(Handle *h, const uint8_t *d_chunk, long s_begin, long s_count);

// f[s] = popcount(row s) / (2 indiv): one warp per SNP row of the staged chunk
__global__ void __launch_bounds__(256)
bed_freq_kernel(const uint8_t *__restrict__ chunk, long rows, long bs, double denom,
                double *__restrict__ f) {
  const long r = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const uint8_t *row = chunk + r * bs;
  unsigned int acc = 0;
  // byte head up to 4-byte alignment, word body, byte tail
  const long head = (4 - (reinterpret_cast<uintptr_t>(row) & 3)) & 3;
  const long h0 = head < bs ? head : bs;
  for (long b = lane; b < h0; b += 32) acc += __popc((unsigned int)row[b]);
  const long nwords = (bs - h0) / 4;
  const uint32_t *w = reinterpret_cast<const uint32_t *>(row + h0);
  for (long i = lane; i < nwords; i += 32) acc += __popc(w[i]);
  for (long b = h0 + nwords * 4 + lane; b < bs; b += 32) acc += __popc((unsigned int)row[b]);
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) f[r] = (double)acc / denom;
}

static long count_newlines(const std::string &path) {
  FILE *fp = fopen(path.c_str(), "rb");
  if (!fp) return -1;
  static char buf[1 << 16];
  long n = 0;
  size_t got;
  while ((got = fread(buf, 1, sizeof(buf), fp)) > 0)
    for (size_t i = 0; i < got; i++) n += buf[i] == '\n';
  fclose(fp);
  return n;
}

// Validation that needs no GPU: suffix, companions, magic, exact size.  Returns 0 and the shape.
static int bed_inspect(const char *path, long *snps, long *indiv, long *bs) {
  if (!path) return fail(-2, __func__, __LINE__, "path is NULL");
  const std::string bed(path);
  if (bed.size() < 4 || bed.compare(bed.size() - 4, 4, ".bed") != 0)
    return fail(-20, __func__, __LINE__, "File not in .bed format");            // read_plink.jl:163-165
  const std::string stem = bed.substr(0, bed.size() - 4);
  const long n_indiv = count_newlines(stem + ".fam"), n_snps = count_newlines(stem + ".bim");
  if (n_indiv < 0 || n_snps < 0)
    return fail(-21, __func__, __LINE__, "Missing supplementary file .fam or .bim");  // :166-168
  FILE *fp = fopen(path, "rb");
  if (!fp) return fail(-22, __func__, __LINE__, "cannot open %s", path);
  unsigned char magic[3] = {0, 0, 0};
  const size_t got = fread(magic, 1, 3, fp);
  fclose(fp);
  if (got != 3 || magic[0] != 0x6c || magic[1] != 0x1b || magic[2] != 0x01)
    return fail(-23, __func__, __LINE__, "Not a correct .bed file");            // :175-177
  if (n_snps <= 0 || n_indiv <= 0)
    return fail(-24, __func__, __LINE__, "empty .bim (%ld lines) or .fam (%ld lines)", n_snps, n_indiv);
  struct stat st;
  if (stat(path, &st) != 0) return fail(-22, __func__, __LINE__, "cannot stat %s", path);
  const long bytes_per_snp = (n_indiv + 3) / 4;
  const long want = 3 + n_snps * bytes_per_snp;
  if ((long)st.st_size < want)
    return fail(-25, __func__, __LINE__, "Too small .bed file (%ld bytes, %ld SNPs x %ld bytes expected)",
                (long)st.st_size, n_snps, bytes_per_snp);                       // unsafe_read throws, :190
  if ((long)st.st_size > want)
    return fail(-26, __func__, __LINE__, "Too large .bed file");                 // :193
  *snps = n_snps;
  *indiv = n_indiv;
  *bs = bytes_per_snp;
  return 0;
}

// Stream the payload: pinned buffer b is filled by fread while buffer b^1 is in flight to the
// device; staged chunk -> bed_freq_kernel + repack_chunk on the handle's stream.
static int bed_stream(Handle *h, const char *path, long bs, long row_offset) {
  long rows_per_chunk = (64L << 20) / bs / 512 * 512;  // ~64 MiB, whole 512-SNP tiles
  if (rows_per_chunk < 512) rows_per_chunk = 512;
  if (rows_per_chunk > round_up(h->snps, 512)) rows_per_chunk = round_up(h->snps, 512);
  const size_t chunk_bytes = (size_t)rows_per_chunk * bs;
  uint8_t *host[2] = {nullptr, nullptr}, *dev[2] = {nullptr, nullptr};
  cudaEvent_t done[2] = {nullptr, nullptr};
  FILE *fp = fopen(path, "rb");
  if (!fp) return fail(-22, __func__, __LINE__, "cannot open %s", path);
  int rc = 0;
  cudaError_t e = cudaSuccess;
  for (int i = 0; i < 2 && e == cudaSuccess; i++) {
    e = cudaMallocHost(&host[i], chunk_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&dev[i], chunk_bytes);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) rc = fail((int)e, __func__, __LINE__, "staging buffers: %s", cudaGetErrorString(e));
  if (!rc && fseek(fp, 3 + row_offset * bs, SEEK_SET) != 0) rc = fail(-22, __func__, __LINE__, "seek failed");
  int b = 0;
  for (long s = 0; s < h->snps && !rc; s += rows_per_chunk, b ^= 1) {
    const long cnt = h->snps - s < rows_per_chunk ? h->snps - s : rows_per_chunk;
    cudaEventSynchronize(done[b]);  // the copy that last used this buffer pair has finished
    if (fread(host[b], 1, (size_t)cnt * bs, fp) != (size_t)cnt * bs) {
      rc = fail(-25, __func__, __LINE__, "short read at SNP %ld", s);
      break;
    }
    e = cudaMemcpyAsync(dev[b], host[b], (size_t)cnt * bs, cudaMemcpyHostToDevice, h->stream);
    if (e != cudaSuccess) { rc = fail((int)e, __func__, __LINE__, "H2D: %s", cudaGetErrorString(e)); break; }
    bed_freq_kernel<<<(unsigned)ceil_div(cnt, 8), 256, 0, h->stream>>>(dev[b], cnt, bs,
                                                                       2.0 * (double)h->indiv, h->d_f + s);
    count_launch();
    rc = repack_chunk(h, dev[b], s, cnt);
    cudaEventRecord(done[b], h->stream);
  }
  cudaStreamSynchronize(h->stream);
  if (!rc && (e = cudaGetLastError()) != cudaSuccess)
    rc = fail((int)e, __func__, __LINE__, "%s", cudaGetErrorString(e));
  fclose(fp);
  for (int i = 0; i < 2; i++) {
    if (host[i]) cudaFreeHost(host[i]);
    if (dev[i]) cudaFree(dev[i]);
    if (done[i]) cudaEventDestroy(done[i]);
  }
  return rc;
}

// SNP rows [row_offset, row_offset + h->snps) of the file into the handle: stream + repack +
// popcount frequencies, then the host copy of f and the row-sum maxima.
int read_bed_rows(Handle *h, const char *path, long row_offset) {
  MB_CHECK(bed_stream(h, path, ceil_div(h->indiv, 4), row_offset));
  h->h_f.resize(h->snps);
  MB_CUDA(cudaMemcpyAsync(h->h_f.data(), h->d_f, sizeof(double) * h->snps, cudaMemcpyDeviceToHost, h->stream));
  MB_CUDA(cudaStreamSynchronize(h->stream));
  return rowsum_maxima(h);
}

}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_bed_shape(const char *bed_path, long *snps, long *indiv) {
  clear_status();
  long s = 0, n = 0, bs = 0;
  MB_CHECK(bed_inspect(bed_path, &s, &n, &bs));
  if (snps) *snps = s;
  if (indiv) *indiv = n;
  return 0;
}

int mb200_read_bed(const char *bed_path, int max_n, void **compressed, long *snps, long *indiv) {
  MB_RANGE();
  clear_status();
  if (!compressed) return fail(-2, __func__, __LINE__, "compressed == NULL");
  *compressed = nullptr;
  long s = 0, n = 0, bs = 0;
  MB_CHECK(bed_inspect(bed_path, &s, &n, &bs));  // before any device work: errors need no GPU
  std::vector<int> devs;
  MB_CHECK(multi_device_list(&devs));
  if (devs.size() > 1) {  // every device streams its own SNP range of the file
    MultiInput in;
    in.src = MultiSource::BedFile;
    in.bed_path = bed_path;
    MB_CHECK(multi_create(devs, s, n, max_n, in, compressed));
  } else {
    Handle *h = nullptr;
    MB_CHECK(create_handle(s, n, max_n, &h, devs.empty() ? -1 : devs[0]));
    const int rc = read_bed_rows(h, bed_path, 0);
    if (rc) { destroy_handle(h); return rc; }
    *compressed = h;
  }
  if (snps) *snps = s;
  if (indiv) *indiv = n;
  return 0;
}

}  // extern "C"
