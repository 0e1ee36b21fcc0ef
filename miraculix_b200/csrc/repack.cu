// repack.cu -- plink2compressed on the GPU.
//
// Replaces plink2gpu (reference src/cuda/dgemm_compressed_cuda.cu:43-170), which uploads TWO
// host-prepared orientations of the bed, and the caller-side host transposes that feed it
// (src/bindings/Julia/compressed_operations.jl:45-66, src/bindings/Fortran/modplink_miraculix.f90:177-218).
// Here only the SNP-major bed is read; one kernel writes both device orientations, recodes
// PLINK codes to genotype values (MXinfo.h:143: 00->0, 01->0, 10->1, 11->2), interleaves the
// fields inside every 32-bit word (see handle.h), zero-pads every row to 128 bytes and
// accumulates the per-SNP allele counts.
//
// HBM-bound: algorithmic bytes = snps*ceil(indiv/4) read + snps*pitch_s + indiv*pitch_i written.
#include <stdlib.h>
#include "common.cuh"
#include "handle.h"

namespace mb200 {

constexpr int RP_TILE = 256;         // SNPs per tile = individuals per tile
constexpr int RP_BYTES = 64;         // input bytes per SNP row of a tile (256 individuals)
constexpr int RP_PITCH = 68;         // smem row pitch (bytes)

// PLINK codes -> values on a whole 32-bit word (16 fields), MXinfo.h:143.
__device__ __forceinline__ uint32_t plink_to_value_word(uint32_t w) {
  // per 2-bit field: value = hi + (hi & lo)  ->  value_hi = hi & lo, value_lo = hi & ~lo
  uint32_t hi = (w >> 1) & 0x55555555u, lo = w & 0x55555555u;
  return ((hi & lo) << 1) | (hi & ~lo);
}
// In-word interleave of the device layout (handle.h): plain field t = 4j + b (bits 8j + 2b)
// moves to bits 8b + 2j, i.e. a 4x4 transpose of 2-bit fields = two delta swaps.
__device__ __forceinline__ uint32_t interleave_word(uint32_t x) {
  uint32_t t = ((x >> 6) ^ x) & 0x00CC00CCu;
  x ^= t ^ (t << 6);
  t = ((x >> 12) ^ x) & 0x0000F0F0u;
  x ^= t ^ (t << 12);
  return x;
}

// plink: rows [s_begin, s_begin + s_count) of the bed, row pitch bs bytes, base = row s_begin.
__global__ void __launch_bounds__(256)
repack_kernel(const uint8_t *__restrict__ plink, long s_begin, long s_count, long snps, long indiv,
              long bs, uint8_t *__restrict__ snp_major, long pitch_s,
              uint8_t *__restrict__ indiv_major, long pitch_i,
              unsigned long long *__restrict__ counts) {
  __shared__ __align__(16) uint8_t tile[RP_TILE * RP_PITCH];  // value-coded, PLAIN field order
  const int tid = threadIdx.x;
  const long i0 = (long)blockIdx.x * RP_TILE;             // first individual of the tile
  const long sl0 = (long)blockIdx.y * RP_TILE;            // first SNP of the tile, chunk-local
  const long w0 = i0 >> 4;                                // first 32-bit word of the tile in a row

  // phase 1: read 256 rows x 16 words, recode, write SNP-major (interleaved), stage in smem
  for (int e = tid; e < RP_TILE * (RP_BYTES / 4); e += 256) {
    const int r = e >> 4, cw = e & 15;
    const long sl = sl0 + r, gw = w0 + cw;
    uint32_t v = 0;
    if (sl < s_count && 4 * gw < bs) {
      const uint8_t *src = plink + sl * bs + 4 * gw;
      const long nbytes = bs - 4 * gw;                    // bytes of this row left (>= 1)
      uint32_t w = src[0];
      if (nbytes > 1) w |= (uint32_t)src[1] << 8;
      if (nbytes > 2) w |= (uint32_t)src[2] << 16;
      if (nbytes > 3) w |= (uint32_t)src[3] << 24;
      v = plink_to_value_word(w);
      const long keep = indiv - 16 * gw;                  // valid fields in this word
      if (keep < 16) v &= (1u << (2 * keep)) - 1u;
      *reinterpret_cast<uint32_t *>(snp_major + (s_begin + sl) * pitch_s + 4 * gw) =
          interleave_word(v);
    }
    *reinterpret_cast<uint32_t *>(tile + r * RP_PITCH + 4 * cw) = v;
    // allele count: a half-warp covers the 16 words of ONE row
    int cnt = __popc(v & 0x55555555u) + 2 * __popc(v & 0xAAAAAAAAu);
    for (int o = 8; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((tid & 15) == 0 && cnt && sl < s_count)
      atomicAdd(&counts[s_begin + sl], (unsigned long long)cnt);
  }
  __syncthreads();

  // phase 2: one interleaved output word = field `il` of 16 consecutive SNP rows
  const long q0 = (s_begin + sl0) >> 4;                   // s_begin, sl0 are multiples of 256
  for (int e = tid; e < RP_TILE * (RP_BYTES / 4); e += 256) {
    const int il = e >> 4, q = e & 15;
    const long i = i0 + il, gq = q0 + q;
    if (i < indiv && 16 * gq < snps) {
      const int cb = il >> 2, sh = 2 * (il & 3);
      const uint8_t *t = tile + (16 * q) * RP_PITCH + cb;
      uint32_t out = 0;
#pragma unroll
      for (int j = 0; j < 4; j++)
#pragma unroll
        for (int b = 0; b < 4; b++)                       // SNP 16q + 4j + b -> bits 8b + 2j
          out |= ((uint32_t)(t[(4 * j + b) * RP_PITCH] >> sh) & 3u) << (8 * b + 2 * j);
      *reinterpret_cast<uint32_t *>(indiv_major + i * pitch_i + 4 * gq) = out;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Fast path (bed rows are whole, 4-byte aligned words: ceil(indiv/4) % 4 == 0): 512 SNP x 512
// individual tiles, so every global access is a full 128-byte line per row -- in (bed row
// segment), out SNP-major (same segment, recoded) and out individual-major (32 words = 512 SNPs of
// one individual).  The tile is transposed IN PLACE in shared memory as 32 x 32 blocks of 16 x 16
// 2-bit fields: a thread loads block (q, cw) (16 words: SNP rows 16q.., word column cw) into
// registers, transposes it with 4 rounds of delta swaps between the words, and stores it as block
// (cw, q).  Smem layout S[k][q][w] (row = 16 q + k, pitch 33 words) makes the block loads (lanes =
// cw), the swapped stores (lanes = cw -> stride 33) and the row-wise copy-out conflict-free.
// ---------------------------------------------------------------------------------------------
constexpr int RF_TILE = 512;
constexpr int RF_THREADS = 512;
constexpr int RF_QP = 33;                       // words per (k, q) row of S
constexpr int RF_KP = 32 * RF_QP + 1;           // words per k plane (+1: 4 rows of a warp hit 4 bank groups)
constexpr int RF_SMEM = 16 * RF_KP * 4 + RF_TILE * 4;   // 67648 bytes of tile + 2 KB of row counts
static_assert(RF_THREADS == RF_TILE, "one row count per thread");

__device__ __forceinline__ void transpose16x16_2bit(uint32_t (&a)[16]) {
  // rounds 1 and 2 exchange whole bytes: one PRMT per output word instead of a 5-instruction
  // delta swap (the kernel is issue/ALU-pipe bound, not HBM bound, once the loads are batched)
#pragma unroll
  for (int k = 0; k < 8; k++) {   // fields 8..15 of word k <-> fields 0..7 of word k + 8
    const uint32_t x = a[k], y = a[k + 8];
    a[k] = __byte_perm(x, y, 0x5410);
    a[k + 8] = __byte_perm(x, y, 0x7632);
  }
#pragma unroll
  for (int k = 0; k < 16; k++)
    if (!(k & 4)) {               // bytes 1,3 of word k <-> bytes 0,2 of word k + 4
      const uint32_t x = a[k], y = a[k + 4];
      a[k] = __byte_perm(x, y, 0x6240);
      a[k + 4] = __byte_perm(x, y, 0x7351);
    }
#pragma unroll
  for (int k = 0; k < 16; k++)
    if (!(k & 2)) { const uint32_t t = ((a[k] >> 4) ^ a[k + 2]) & 0x0F0F0F0Fu; a[k + 2] ^= t; a[k] ^= t << 4; }
#pragma unroll
  for (int k = 0; k < 16; k++)
    if (!(k & 1)) { const uint32_t t = ((a[k] >> 2) ^ a[k + 1]) & 0x33333333u; a[k + 1] ^= t; a[k] ^= t << 2; }
}

template <int VW> struct RfVec;
template <> struct RfVec<1> { typedef uint32_t T; };
template <> struct RfVec<2> { typedef uint2 T; };
template <> struct RfVec<4> { typedef uint4 T; };

// Phase 1 of the fast kernel: 32 / VW lanes per SNP row; read, recode, count, write SNP-major, stage
// plain (un-interleaved) in smem.  The bed loads of a whole batch of rows are issued before the first
// one is used: with one load in flight per thread the phase was a chain of NIT DRAM round trips
// (ncu: 46 % of all stall samples on the first use of the loaded word).  After that fix the phase is
// issue-bound, so everything per row is strength-reduced: running pointers instead of 64-bit
// row * pitch products, constant smem offsets, row counts parked in smem (one coalesced atomic per
// row after the barrier), and FULL = interior tile (no row / column / ragged-word predicates).
template <int VW, bool FULL>
__device__ __forceinline__ void rf_phase1(const uint32_t *__restrict__ plink, long s_begin, long s_count,
                                          long indiv, long wpr, uint8_t *__restrict__ snp_major,
                                          long pitch_s, long w0, long sl0, uint32_t *S, int *scnt,
                                          int tid) {
  typedef typename RfVec<VW>::T Vec;
  constexpr int LPR = 32 / VW;                          // lanes per row
  constexpr int RPI = RF_THREADS / LPR;                 // rows per iteration: 16 / 32 / 64
  constexpr int NIT = RF_TILE / RPI;                    // iterations: 32 / 16 / 8
  constexpr int BATCH = VW == 4 ? 4 : 8;                // loads in flight per thread (<= 16 registers)
  constexpr int DST_STEP = (RPI / 16) * RF_QP;          // (r0 + i RPI) & 15 == r0 & 15, >> 4 grows by RPI / 16
  const int wl = (tid % LPR) * VW;                      // first word of this thread in the segment
  const long gw = w0 + wl;
  const int r0 = tid / LPR;
  const bool in_cols = FULL || gw < wpr;                // wpr % VW == 0: the vector is whole
  const Vec *src = reinterpret_cast<const Vec *>(plink + (sl0 + r0) * wpr + gw);
  const long src_step = (long)RPI * wpr / VW;
  uint8_t *out = snp_major + (s_begin + sl0 + r0) * pitch_s + 4 * gw;
  const long out_step = (long)RPI * pitch_s;
  uint32_t *dst = S + (r0 & 15) * RF_KP + (r0 >> 4) * RF_QP + wl;
  int *cdst = scnt + r0;
  long rows_left = s_count - (sl0 + r0);                // row r of this thread exists iff rows_left > 0
#pragma unroll 1
  for (int it0 = 0; it0 < NIT; it0 += BATCH) {
    union { Vec vec; uint32_t w[VW]; } raw[BATCH];      // Vec-aligned wherever the compiler keeps it
#pragma unroll
    for (int b = 0; b < BATCH; b++) {
      if (FULL || (in_cols && rows_left > (long)b * RPI)) {
        raw[b].vec = __ldg(src);
      } else {
#pragma unroll
        for (int j = 0; j < VW; j++) raw[b].w[j] = 0;   // PLINK code 00 -> value 0
      }
      src += src_step;
    }
#pragma unroll
    for (int b = 0; b < BATCH; b++) {
      uint32_t v[VW];
      union { Vec vec; uint32_t w[VW]; } o;
      int cnt = 0;
#pragma unroll
      for (int j = 0; j < VW; j++) {
        v[j] = plink_to_value_word(raw[b].w[j]);
        if (!FULL) {
          const long keep = indiv - 16 * (gw + j);      // valid fields in this word
          if (keep < 16) v[j] = keep > 0 ? v[j] & ((1u << (2 * keep)) - 1u) : 0u;
        }
        cnt += __popc(v[j]) + __popc(v[j] & 0xAAAAAAAAu);   // sum of 2-bit values
        o.w[j] = interleave_word(v[j]);
      }
      if (FULL || rows_left > (long)b * RPI)            // data, or row padding up to the 128-byte pitch
        *reinterpret_cast<Vec *>(out) = o.vec;
      out += out_step;
#pragma unroll
      for (int j = 0; j < VW; j++) dst[b * DST_STEP + j] = v[j];
#pragma unroll
      for (int o2 = LPR / 2; o2; o2 >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o2);
      if ((tid % LPR) == 0) cdst[b * RPI] = cnt;
    }
    dst += BATCH * DST_STEP;
    cdst += BATCH * RPI;
    rows_left -= (long)BATCH * RPI;
  }
}

// VW = words per thread in phase 1 (1, 2 or 4: what the alignment of the bed rows allows)
template <int VW>
__global__ void __launch_bounds__(RF_THREADS, 3)
repack_fast_kernel(const uint32_t *__restrict__ plink, long s_begin, long s_count, long snps,
                   long indiv, long wpr /* words per bed row */, uint8_t *__restrict__ snp_major,
                   long pitch_s, uint8_t *__restrict__ indiv_major, long pitch_i,
                   unsigned long long *__restrict__ counts) {
  extern __shared__ uint32_t S[];
  int *scnt = reinterpret_cast<int *>(S + 16 * RF_KP);  // allele count of every SNP row of the tile
  const int tid = threadIdx.x, lane = tid & 31;
  const long w0 = (long)blockIdx.x * 32;                  // first word column (16 individuals each)
  const long sl0 = (long)blockIdx.y * RF_TILE;            // first SNP of the tile, chunk-local
  const bool edge = 16 * (w0 + 32) > indiv;               // tile holds the ragged end of the rows

  if (!edge && sl0 + RF_TILE <= s_count)
    rf_phase1<VW, true>(plink, s_begin, s_count, indiv, wpr, snp_major, pitch_s, w0, sl0, S, scnt, tid);
  else
    rf_phase1<VW, false>(plink, s_begin, s_count, indiv, wpr, snp_major, pitch_s, w0, sl0, S, scnt, tid);
  __syncthreads();
  {                                                       // RF_THREADS == RF_TILE: one row per thread
    const int c = scnt[tid];
    if (c && sl0 + tid < s_count) atomicAdd(&counts[s_begin + sl0 + tid], (unsigned long long)c);
  }

  // phase 2: in-place transpose, two blocks per thread (block id = q * 32 + cw, lanes = cw).
  // SNP row k of a block is loaded into register 4 (k % 4) + k / 4, so the transposed word comes
  // out with its 16 SNP fields already in the interleaved order of the device layout.
  uint32_t a0[16], a1[16];
  const int q0b = tid >> 5, q1b = q0b + 16, cw = lane;
#pragma unroll
  for (int k = 0; k < 16; k++) {
    a0[4 * (k & 3) + (k >> 2)] = S[k * RF_KP + q0b * RF_QP + cw];
    a1[4 * (k & 3) + (k >> 2)] = S[k * RF_KP + q1b * RF_QP + cw];
  }
  __syncthreads();
  transpose16x16_2bit(a0);
  transpose16x16_2bit(a1);
#pragma unroll
  for (int il = 0; il < 16; il++) {   // word il = individual 16 cw + il, fields = SNPs 16 q ..
    S[il * RF_KP + cw * RF_QP + q0b] = a0[il];
    S[il * RF_KP + cw * RF_QP + q1b] = a1[il];
  }
  __syncthreads();

  // phase 3: 8 lanes per individual row of the tile (4 words = 16 bytes each), 128 B per row
  const long gq0 = (s_begin + sl0) >> 4;                  // s_begin, sl0 are multiples of 512
  const int wl = (tid & 7) * 4;
  const long gq = gq0 + wl;
  const long qwords = pitch_i >> 2;                       // whole pitch: SNP rows past the end are 0
#pragma unroll 2
  for (int r = tid >> 3; r < RF_TILE; r += RF_THREADS / 8) {
    const long i = 16 * w0 + r;
    if (i < indiv && gq < qwords) {                       // the kernel writes the row padding too
      const uint32_t *src = S + (r & 15) * RF_KP + (r >> 4) * RF_QP + wl;
      *reinterpret_cast<uint4 *>(indiv_major + i * pitch_i + 4 * gq) =
          make_uint4(src[0], src[1], src[2], src[3]);
    }
  }
}

int repack_chunk(Handle *h, const uint8_t *d_chunk, long s_begin, long s_count) {
  const long bs = ceil_div(h->indiv, 4);
  static const int use_fast = [] {
    const char *e = getenv("MB200_REPACK_FAST");
    return e ? atoi(e) : 1;
  }();
  {  // the > 48 KB dynamic shared memory opt-in is per device
    static std::atomic<unsigned long long> done{0};
    if (h->device < 0 || h->device >= 64 || !(done.load(std::memory_order_acquire) >> h->device & 1ull)) {
      MB_CUDA(cudaFuncSetAttribute(repack_fast_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, RF_SMEM));
      MB_CUDA(cudaFuncSetAttribute(repack_fast_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, RF_SMEM));
      MB_CUDA(cudaFuncSetAttribute(repack_fast_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, RF_SMEM));
      if (h->device >= 0 && h->device < 64) done.fetch_or(1ull << h->device, std::memory_order_release);
    }
  }
  if (use_fast && bs % 4 == 0 && (reinterpret_cast<uintptr_t>(d_chunk) & 3) == 0 &&
      s_begin % RF_TILE == 0) {
    dim3 grid((unsigned)ceil_div(bs / 4, 32), (unsigned)ceil_div(s_count, RF_TILE));
    const uint32_t *src = reinterpret_cast<const uint32_t *>(d_chunk);
    const uintptr_t al = reinterpret_cast<uintptr_t>(d_chunk);
#define RF_LAUNCH(VW)                                                                          \
  repack_fast_kernel<VW><<<grid, RF_THREADS, RF_SMEM, h->stream>>>(                              \
      src, s_begin, s_count, h->snps, h->indiv, bs / 4, h->d_snp_major, h->pitch_s,             \
      h->d_indiv_major, h->pitch_i, h->d_counts)
    if (bs % 16 == 0 && al % 16 == 0) RF_LAUNCH(4);
    else if (bs % 8 == 0 && al % 8 == 0) RF_LAUNCH(2);
    else RF_LAUNCH(1);
#undef RF_LAUNCH
    count_launch();
    MB_LAUNCH_CHECK();
    return 0;
  }
  if (!h->zeroed) {  // the generic kernel relies on zero-initialised padding
    MB_CUDA(cudaMemsetAsync(h->d_snp_major, 0, (size_t)h->snps * h->pitch_s, h->stream));
    MB_CUDA(cudaMemsetAsync(h->d_indiv_major, 0, (size_t)h->indiv * h->pitch_i, h->stream));
    h->zeroed = true;
  }
  dim3 grid((unsigned)ceil_div(h->indiv, RP_TILE), (unsigned)ceil_div(s_count, RP_TILE));
  repack_kernel<<<grid, 256, 0, h->stream>>>(d_chunk, s_begin, s_count, h->snps, h->indiv, bs,
                                             h->d_snp_major, h->pitch_s, h->d_indiv_major,
                                             h->pitch_i, h->d_counts);
  count_launch();
  MB_LAUNCH_CHECK();
  return 0;
}

int repack_from_device(Handle *h, const uint8_t *d_plink) {
  // grid.y limit 65535 tiles = 16.7M SNPs per launch; chunk to stay below
  const long bs = ceil_div(h->indiv, 4);
  const long max_rows = 65534L * RP_TILE;  // multiple of 512: keeps the fast path's tile alignment
  for (long s = 0; s < h->snps; s += max_rows) {
    long cnt = h->snps - s < max_rows ? h->snps - s : max_rows;
    MB_CHECK(repack_chunk(h, d_plink + s * bs, s, cnt));
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// freq = counts / (2 indiv)  (popcount rule, reference src/bindings/Julia/read_plink.jl:197-203)
// ---------------------------------------------------------------------------------------------
__global__ void freq_kernel(const unsigned long long *__restrict__ counts, long snps, double denom,
                            double *__restrict__ f) {
  long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < snps) f[j] = (double)counts[j] / denom;
}
int freq_from_counts(Handle *h) {
  freq_kernel<<<(unsigned)ceil_div(h->snps, 256), 256, 0, h->stream>>>(
      h->d_counts, h->snps, 2.0 * (double)h->indiv, h->d_f);
  count_launch();
  MB_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// max_r sum_k z_rk of one device orientation: the rigorous error bound of the digit-sliced
// tensor-core dgemm (dgemm_tc.cu) needs it.  One warp per row, HBM-bound, once per handle.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rowsum_max_kernel(const uint8_t *__restrict__ A, long pitch, long rows, unsigned int *__restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long warp0 = (long)blockIdx.x * 8 + (threadIdx.x >> 5), nwarps = (long)gridDim.x * 8;
  unsigned int best = 0;
  for (long r = warp0; r < rows; r += nwarps) {
    const uint4 *row = reinterpret_cast<const uint4 *>(A + r * pitch);
    unsigned int acc = 0;
    for (long i = lane; i < pitch / 16; i += 32) {
      const uint4 v = __ldg(row + i);
      acc += __popc(v.x & 0x55555555u) + 2 * __popc(v.x & 0xAAAAAAAAu) +
             __popc(v.y & 0x55555555u) + 2 * __popc(v.y & 0xAAAAAAAAu) +
             __popc(v.z & 0x55555555u) + 2 * __popc(v.z & 0xAAAAAAAAu) +
             __popc(v.w & 0x55555555u) + 2 * __popc(v.w & 0xAAAAAAAAu);
    }
    acc = __reduce_add_sync(0xffffffffu, acc);
    best = acc > best ? acc : best;
  }
  if (lane == 0 && best) atomicMax(out, best);
}
// The SNP-major row sums ARE the allele counts the repack kernels accumulate (after the missing ->
// 0 recode and the edge mask), so that orientation's maximum is a reduction over 8 bytes per SNP
// instead of a second read of the whole packed matrix.
__global__ void __launch_bounds__(256)
counts_max_kernel(const unsigned long long *__restrict__ counts, long snps, unsigned int *__restrict__ out) {
  unsigned long long best = 0;
  for (long j = (long)blockIdx.x * blockDim.x + threadIdx.x; j < snps; j += (long)gridDim.x * blockDim.x) {
    const unsigned long long c = __ldg(counts + j);
    best = c > best ? c : best;
  }
  unsigned int b32 = (unsigned int)best;                  // <= 2 * indiv < 2^32 (indiv is an int)
  b32 = __reduce_max_sync(0xffffffffu, b32);
  if ((threadIdx.x & 31) == 0 && b32) atomicMax(out, b32);
}
int rowsum_maxima(Handle *h) {
  unsigned int *d = reinterpret_cast<unsigned int *>(h->d_w);  // scratch, free at creation time
  unsigned int host[2] = {0, 0};
  MB_CUDA(cudaMemsetAsync(d, 0, 2 * sizeof(unsigned int), h->stream));
  rowsum_max_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(h->d_indiv_major, h->pitch_i, h->indiv, d);
  counts_max_kernel<<<h->sm_count * 4, 256, 0, h->stream>>>(h->d_counts, h->snps, d + 1);
  count_launch(2);
  MB_LAUNCH_CHECK();
  MB_CUDA(cudaMemcpyAsync(host, d, sizeof(host), cudaMemcpyDeviceToHost, h->stream));
  MB_CUDA(cudaStreamSynchronize(h->stream));
  h->rowsum_max[0] = host[0];
  h->rowsum_max[1] = host[1];
  return 0;
}

// ---------------------------------------------------------------------------------------------
// On-device simulator (benchmark inputs): p_j ~ U(0.1,0.6),
// z_ij ~ Binomial(2,p_j) = [u1<p] + [u2<p], PLINK codes 0->00, 1->10, 2->11, pad bits 00.
// Counter-based: every genotype is a pure function of (seed, snp, indiv).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__global__ void __launch_bounds__(256)
simulate_kernel(uint8_t *__restrict__ plink, long snps, long indiv, long bs,
                unsigned long long seed, long s_offset) {
  const long total_words = snps * ((bs + 3) / 4);
  const long wpr = (bs + 3) / 4;  // 32-bit words per row (last one partial)
  for (long w = (long)blockIdx.x * blockDim.x + threadIdx.x; w < total_words;
       w += (long)gridDim.x * blockDim.x) {
    const long s = w / wpr, wb = w % wpr;
    const uint64_t ks = mix64(seed ^ (0xA5A5A5A5ull + (uint64_t)(s + s_offset) * 0x100000001B3ull));
    const uint32_t thr = (uint32_t)((0.1 + 0.5 * ((ks >> 11) * (1.0 / 9007199254740992.0))) *
                                    4294967296.0);
    uint32_t word = 0;
#pragma unroll
    for (int g = 0; g < 16; g += 2) {  // one 64-bit hash -> two genotypes
      const long i = wb * 16 + g;
      uint64_t r = mix64(ks + (uint64_t)i * 0xD6E8FEB86659FD93ull);
      uint32_t a0 = (uint32_t)r < thr, a1 = (uint32_t)(r >> 32) < thr;
      r = mix64(r);
      uint32_t c0 = (uint32_t)r < thr, c1 = (uint32_t)(r >> 32) < thr;
      uint32_t z0 = a0 + a1, z1 = c0 + c1;
      uint32_t code0 = i < indiv ? (z0 == 0 ? 0u : z0 + 1u) : 0u;
      uint32_t code1 = (i + 1) < indiv ? (z1 == 0 ? 0u : z1 + 1u) : 0u;
      word |= (code0 << (2 * g)) | (code1 << (2 * g + 2));
    }
    uint8_t *dst = plink + s * bs + wb * 4;
    const long rem = bs - wb * 4;
    if (rem >= 4 && ((s * bs) & 3) == 0) {
      *reinterpret_cast<uint32_t *>(dst) = word;
    } else {
      for (int t = 0; t < 4 && t < rem; t++) dst[t] = (uint8_t)(word >> (8 * t));
    }
  }
}
int simulate_plink(uint8_t *d_plink, long snps, long indiv, unsigned long long seed,
                   cudaStream_t stream, long s_offset) {
  const long bs = ceil_div(indiv, 4);
  int dev = 0;
  cudaGetDevice(&dev);
  simulate_kernel<<<device_sm_count(dev) * 8, 256, 0, stream>>>(d_plink, snps, indiv, bs, seed, s_offset);
  count_launch();
  MB_LAUNCH_CHECK();
  return 0;
}

}  // namespace mb200
