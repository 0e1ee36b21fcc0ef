// crossprod_kernels.cuh -- small HBM-bound kernels around the crossproduct GEMM
// (SURVEY.md 8a rows a9, a11, a12, a13).
#pragma once
#include <stdlib.h>
#include "common.cuh"

namespace mb200 {

// PLINK -> 2bit byte table, generated on the host at first use by the rule of the reference's
// generator (src/bindings/Julia/read_plink.jl:50-80): most significant pair first, "10" -> 1,
// "11" -> 2, any "01" pair turns the whole byte into 0xff.  Equals d_conversion_table
// (src/cuda/snp_multiply_cuda.h:202); tests/test_golden.py checks all 256 entries.
__constant__ uint8_t c_plink2twobit[256];

inline void host_plink2twobit(uint8_t *table) {
  for (int v = 0; v < 256; v++) {
    uint8_t out = 0;
    for (int pos = 3; pos >= 0; pos--) {
      const int c = (v >> (2 * pos)) & 3;
      out = (uint8_t)(out << 2);
      if (c == 2) out |= 1;
      else if (c == 3) out |= 2;
      else if (c == 1) { out = 0xff; break; }
    }
    table[v] = out;
  }
}

// Expand K-slab [kb0, kb0 + kbytes) (in packed bytes) of the 2-bit matrix x[rows][nb] to int8
// X8[rows_pad][kpad] (kpad = 4*kbytes rounded up to 128); every 2-bit field is the value 0..3
// (replaces the two masked u4 passes of src/cuda/snp_multiply_cuda.h:170-192 and, with
// is_plink, device_convert_plink_2bit :370-373).  One thread per 4 packed bytes -> 16 B store.
// Rows >= rows and columns >= 4*kbytes are written as zero.
__global__ void __launch_bounds__(256)
expand_2bit_kernel(const uint8_t *__restrict__ x, long rows, long nb, long kb0, long kbytes,
                   int is_plink, int8_t *__restrict__ X8, long rows_pad, long kpad) {
  const long groups_per_row = kpad / 16;
  const long total = rows_pad * groups_per_row;
  for (long g = (long)blockIdx.x * blockDim.x + threadIdx.x; g < total;
       g += (long)gridDim.x * blockDim.x) {
    const long r = g / groups_per_row, gc = g % groups_per_row;
    uint32_t out[4] = {0u, 0u, 0u, 0u};
    if (r < rows) {
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const long b = gc * 4 + t;
        if (b < kbytes) {
          uint32_t v = x[r * nb + kb0 + b];
          if (is_plink) v = c_plink2twobit[v];
          out[t] = (v & 3u) | ((v & 0xCu) << 6) | ((v & 0x30u) << 12) | ((v & 0xC0u) << 18);
        }
      }
    }
    *reinterpret_cast<uint4 *>(X8 + r * kpad + gc * 16) = make_uint4(out[0], out[1], out[2], out[3]);
  }
}

// Word path of the expansion (the code IS the value: not PLINK-coded; rows and slab start whole
// 32-bit words): one 4-byte load and one 16-byte store per thread, four rows in flight, no 64-bit
// divisions -- the byte kernel above was issue bound (68 % issue slots, 44 % of DRAM peak).
// grid.x covers the 16-byte groups of a row, grid.y strides over the rows.
__global__ void __launch_bounds__(256)
expand_2bit_words_kernel(const uint32_t *__restrict__ x, long rows, long nw, long kw0, long kbytes,
                         int8_t *__restrict__ X8, long rows_pad, long kpad) {
  const long gc = (long)blockIdx.x * blockDim.x + threadIdx.x;   // packed word of the slab
  if (gc >= kpad / 16) return;
  const long left = kbytes - 4 * gc;                             // packed bytes of this word inside the slab
  const uint32_t keep = left >= 4 ? 0xFFFFFFFFu : left > 0 ? (1u << (8 * left)) - 1u : 0u;
  const uint32_t *src = x + kw0 + gc;
  uint4 *dst = reinterpret_cast<uint4 *>(X8 + gc * 16);
  const long pitch4 = kpad / 16;
  for (long r0 = (long)blockIdx.y * 4; r0 < rows_pad; r0 += (long)gridDim.y * 4) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; i++) w[i] = (keep && r0 + i < rows) ? __ldg(src + (r0 + i) * nw) & keep : 0u;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      if (r0 + i >= rows_pad) break;
      uint32_t o[4];
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const uint32_t v = __byte_perm(w[i], 0u, 0x4440 + t);    // byte t, zero-extended
        o[t] = (v | (v << 6) | (v << 12) | (v << 18)) & 0x03030303u;
      }
      dst[(r0 + i) * pitch4] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// Launches the expansion of one K-slab (word path when the layout allows it).
inline void expand_2bit_launch(const uint8_t *d_x, long rows, long nb, long kb0, long kbytes,
                               int is_plink, int8_t *X8, long rows_pad, long kpad, int sms,
                               cudaStream_t stream) {
  static const bool bytes_only = getenv("MB200_EXPAND_BYTES") != nullptr;  // A/B switch
  if (!bytes_only && !is_plink && nb % 4 == 0 && kb0 % 4 == 0 && (reinterpret_cast<uintptr_t>(d_x) & 3) == 0) {
    const unsigned gx = (unsigned)ceil_div(kpad / 16, 256);
    long gy = ceil_div((long)sms * 16, (long)gx);                // ~16 blocks per SM in flight
    if (gy > ceil_div(rows_pad, 4)) gy = ceil_div(rows_pad, 4);
    if (gy > 65535) gy = 65535;
    expand_2bit_words_kernel<<<dim3(gx, (unsigned)gy), 256, 0, stream>>>(
        reinterpret_cast<const uint32_t *>(d_x), rows, nb / 4, kb0 / 4, kbytes, X8, rows_pad, kpad);
  } else {
    expand_2bit_kernel<<<sms * 8, 256, 0, stream>>>(d_x, rows, nb, kb0, kbytes, is_plink, X8, rows_pad,
                                                    kpad);
  }
}

// int32 counts -> fp64, mirrored into the full symmetric matrix; with `affine` the GRM update of
// crossproduct.jl:95-106, G = (M - (m_i + m_j)/n + total/n^2) / c, is applied on the way out as
//   G = (n^2 M - n m_i - (n m_j - total)) * (1 / (c n^2))
// with the bracket EXACT in 64-bit integers (M, m and total are integers; n^2 * 4k < 2^62 is
// checked on the host) and one rounding each for the conversion and the scale: at least as
// accurate as the BLAS sequence of the reference, exactly symmetric, and ONE fp64 instruction per
// entry instead of twelve (fp64 instructions in the epilogue warps of the tensor-core GEMM stall
// on the math pipe the MMAs keep busy; they cost a quarter of the GEMM time).
// scal[0] = total = sum m, scal[1] = c = 2 sum f (1 - f).
// (CpAffine / cp_affine live in common.cuh: the tensor-core GEMM epilogue uses them too)
__global__ void __launch_bounds__(256)
counts_to_double_kernel(const int32_t *__restrict__ M32, long ld32, long n,
                        double *__restrict__ ans, int affine, const double *__restrict__ m,
                        const double *__restrict__ scal, int do_scale) {
  __shared__ int32_t t[32][33];
  const long bi = blockIdx.x, bj = blockIdx.y;  // output tile: rows i in bi, cols j in bj
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const CpAffine aff = cp_affine_setup(n, affine, do_scale, scal);
  if (bi <= bj) {
    // source M32[i][j], contiguous in j: load (i = ty.., j = tx), transpose through smem
    for (int yy = ty; yy < 32; yy += 8) {
      const long i = bi * 32 + yy, j = bj * 32 + tx;
      t[yy][tx] = (i < n && j < n) ? M32[(i <= j ? i : j) * ld32 + (i <= j ? j : i)] : 0;
    }
    __syncthreads();
    for (int yy = ty; yy < 32; yy += 8) {
      const long j = bj * 32 + yy, i = bi * 32 + tx;
      if (i < n && j < n) {
        ans[i + j * n] = affine ? cp_affine_m(t[tx][yy], m[i], m[j], aff) : (double)t[tx][yy];
      }
    }
  } else {
    // lower tile: value = M32[j][i], contiguous in i -> direct coalesced read and write
    for (int yy = ty; yy < 32; yy += 8) {
      const long j = bj * 32 + yy, i = bi * 32 + tx;
      if (i < n && j < n) {
        const int cnt = M32[j * ld32 + i];
        ans[i + j * n] = affine ? cp_affine_m(cnt, m[i], m[j], aff) : (double)cnt;
      }
    }
  }
}

// column sums m_j = sum_i M[i,j] (col_sum of src/bindings/Julia/crossproduct.jl:92); one warp
// per column, exact while the sums stay below 2^53 (integer-valued doubles).
__global__ void __launch_bounds__(256)
colsum_kernel(const double *__restrict__ M, long n, double *__restrict__ m) {
  const long j = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (j >= n) return;
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  for (long i = lane; i < n; i += 32) acc += M[i + j * n];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) m[j] = acc;
}
// single-block reductions: total = sum m_j ;  c = 2 sum f (1 - f)
__global__ void __launch_bounds__(1024)
reduce_sum_kernel(const double *__restrict__ v, long n, int mode, double *__restrict__ out) {
  __shared__ double red[32];
  double acc = 0.0;
  for (long i = threadIdx.x; i < n; i += 1024) acc += mode ? v[i] * (1.0 - v[i]) : v[i];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = red[threadIdx.x];
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) *out = mode ? 2.0 * acc : acc;
  }
}
// LD epilogue (crossproduct.jl:128-152): M -= 4 n_obs f_i f_j ; then / sigma_i / sigma_j with
// sigma = sqrt(diag).  Two passes because the diagonal must be final first.
__global__ void __launch_bounds__(256)
ld_rank1_kernel(double *__restrict__ M, long n, const double *__restrict__ f, double alpha,
                double *__restrict__ sigma) {
  const long total = n * n;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long)gridDim.x * blockDim.x) {
    const long i = e % n, j = e / n;
    // symmetric by construction: use the (min,max) ordering of the syr! upper update
    const double fi = f[i < j ? i : j], fj = f[i < j ? j : i];
    const double v = __dadd_rn(M[e], __dmul_rn(__dmul_rn(alpha, fi), fj));
    M[e] = v;
    if (i == j) sigma[i] = sqrt(v);
  }
}
__global__ void __launch_bounds__(256)
ld_scale_kernel(double *__restrict__ M, long n, const double *__restrict__ sigma) {
  const long total = n * n;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long)gridDim.x * blockDim.x) {
    const long i = e % n, j = e / n;
    M[e] = M[e] / sigma[i] / sigma[j];
  }
}

// ---- exact row sums of M = X X^T without forming M:  m = X (X^T 1)  (used by the partitioned
// GRM so that the centring needs no communication between the GPUs that own the output tiles)
__device__ __forceinline__ uint32_t packed_values(uint32_t byte, int is_plink) {
  return is_plink ? c_plink2twobit[byte] : byte;  // 4 fields, each the value 0..3
}
// s[4b + t] += sum_r field_t(x[r][b]);  thread = packed byte column, grid.y splits the rows.
// Sums stay below 3 rows < 2^32.  (Byte path: any nb; the word path below needs nb % 4 == 0.)
__global__ void __launch_bounds__(256)
snp_colsum_kernel(const uint8_t *__restrict__ x, long rows, long nb, int is_plink,
                  unsigned int *__restrict__ s) {
  const long b = (long)blockIdx.x * 256 + threadIdx.x;
  if (b >= nb) return;
  const long per = (rows + gridDim.y - 1) / gridDim.y;
  const long r0 = (long)blockIdx.y * per, r1 = r0 + per < rows ? r0 + per : rows;
  unsigned int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  for (long r = r0; r < r1; r++) {
    const uint32_t v = packed_values(x[r * nb + b], is_plink);
    c0 += v & 3u; c1 += (v >> 2) & 3u; c2 += (v >> 4) & 3u; c3 += (v >> 6) & 3u;
  }
  if (c0) atomicAdd(&s[4 * b + 0], c0);
  if (c1) atomicAdd(&s[4 * b + 1], c1);
  if (c2) atomicAdd(&s[4 * b + 2], c2);
  if (c3) atomicAdd(&s[4 * b + 3], c3);
}
// m[r] = sum_k x_rk s_k, exact in 64-bit integers (one warp per row)
__global__ void __launch_bounds__(256)
row_dot_kernel(const uint8_t *__restrict__ x, long rows, long nb, int is_plink,
               const unsigned int *__restrict__ s, double *__restrict__ m) {
  const long r = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  unsigned long long acc = 0;
  for (long b = lane; b < nb; b += 32) {
    const uint32_t v = packed_values(x[r * nb + b], is_plink);
    acc += (unsigned long long)((v & 3u) * s[4 * b] + ((v >> 2) & 3u) * s[4 * b + 1]) +
           (unsigned long long)(((v >> 4) & 3u) * s[4 * b + 2] + ((v >> 6) & 3u) * s[4 * b + 3]);
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) m[r] = (double)acc;
}
// ---- word variants (rows are whole 32-bit words: nb % 4 == 0, 4-byte aligned base) ----
__device__ __forceinline__ uint32_t word_values(uint32_t w, int is_plink) {
  if (!is_plink) return w;
  return (uint32_t)c_plink2twobit[w & 255u] | ((uint32_t)c_plink2twobit[(w >> 8) & 255u] << 8) |
         ((uint32_t)c_plink2twobit[(w >> 16) & 255u] << 16) | ((uint32_t)c_plink2twobit[w >> 24] << 24);
}
// thread = word column (16 genotypes); the four fields of every byte are counted in four SWAR
// byte-lane accumulators, flushed every 64 rows (64 x 3 < 256).
__global__ void __launch_bounds__(128)
snp_colsum_words_kernel(const uint32_t *__restrict__ x, long rows, long nw, int is_plink,
                        unsigned int *__restrict__ s) {
  const long w = (long)blockIdx.x * 128 + threadIdx.x;
  if (w >= nw) return;
  const long per = (rows + gridDim.y - 1) / gridDim.y;
  const long r0 = (long)blockIdx.y * per, r1 = r0 + per < rows ? r0 + per : rows;
  unsigned int c[16];
#pragma unroll
  for (int i = 0; i < 16; i++) c[i] = 0;
  for (long rb = r0; rb < r1; rb += 64) {
    const long re = rb + 64 < r1 ? rb + 64 : r1;
    uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (long r = rb; r < re; r++) {
      const uint32_t v = word_values(x[r * nw + w], is_plink);
      a0 += v & 0x03030303u;
      a1 += (v >> 2) & 0x03030303u;
      a2 += (v >> 4) & 0x03030303u;
      a3 += (v >> 6) & 0x03030303u;
    }
#pragma unroll
    for (int b = 0; b < 4; b++) {
      c[4 * b + 0] += (a0 >> (8 * b)) & 255u;
      c[4 * b + 1] += (a1 >> (8 * b)) & 255u;
      c[4 * b + 2] += (a2 >> (8 * b)) & 255u;
      c[4 * b + 3] += (a3 >> (8 * b)) & 255u;
    }
  }
#pragma unroll
  for (int i = 0; i < 16; i++)
    if (c[i]) atomicAdd(&s[16 * w + i], c[i]);
}
// Byte planes of the column sums for dp4a: word w, field j (genotypes 16w + 4b + j, b = 0..3,
// i.e. the four byte lanes of (v >> 2j) & 0x03030303) -> planes[(4w + j)] = uint4 whose component
// p packs bits [8p, 8p+8) of those four sums.  In place over s (same size, one thread per (w, j)
// group would race) -> separate output.
__global__ void __launch_bounds__(256)
colsum_planes_kernel(const unsigned int *__restrict__ s, long nw, uint4 *__restrict__ planes) {
  const long g = (long)blockIdx.x * 256 + threadIdx.x;  // g = 4 w + j
  if (g >= 4 * nw) return;
  const long w = g >> 2;
  const int j = (int)(g & 3);
  unsigned int v[4];
#pragma unroll
  for (int b = 0; b < 4; b++) v[b] = s[16 * w + 4 * b + j];
  unsigned int pl[4];
#pragma unroll
  for (int p = 0; p < 4; p++)
    pl[p] = ((v[0] >> (8 * p)) & 255u) | (((v[1] >> (8 * p)) & 255u) << 8) |
            (((v[2] >> (8 * p)) & 255u) << 16) | (((v[3] >> (8 * p)) & 255u) << 24);
  planes[g] = make_uint4(pl[0], pl[1], pl[2], pl[3]);
}
// m[r] = sum_k x_rk s_k with u8 x u8 dot products (dp4a) against the byte planes: one warp per 8
// rows, the planes of a word are loaded once and applied to the 8 rows; each plane accumulates in
// 32 bits (<= 4 x 3 x 255 per dp4a, 4 per word, nw/32 words per lane), combined in 64 bits.
constexpr int RD_ROWS = 8;
__global__ void __launch_bounds__(256)
row_dot_words_kernel(const uint32_t *__restrict__ x, long rows, long nw, int is_plink,
                     const uint4 *__restrict__ planes, double *__restrict__ m) {
  const long r0 = ((long)blockIdx.x * 8 + (threadIdx.x >> 5)) * RD_ROWS;
  if (r0 >= rows) return;
  const int lane = threadIdx.x & 31;
  unsigned int acc[RD_ROWS][4];
#pragma unroll
  for (int q = 0; q < RD_ROWS; q++)
#pragma unroll
    for (int p = 0; p < 4; p++) acc[q][p] = 0;
  unsigned long long tot[RD_ROWS];
#pragma unroll
  for (int q = 0; q < RD_ROWS; q++) tot[q] = 0;
  long it = 0;
  const long wper = (nw + gridDim.y - 1) / gridDim.y;
  const long w1 = (blockIdx.y + 1) * wper < nw ? (blockIdx.y + 1) * wper : nw;
  for (long w = blockIdx.y * wper + lane; w < w1; w += 32, it++) {
    uint4 pl[4];
#pragma unroll
    for (int j = 0; j < 4; j++) pl[j] = planes[4 * w + j];
    uint32_t xv[RD_ROWS];  // all loads first (rows past the end re-read the last row, discarded)
#pragma unroll
    for (int q = 0; q < RD_ROWS; q++) {
      const long rr = r0 + q < rows ? r0 + q : rows - 1;
      xv[q] = x[rr * nw + w];
    }
#pragma unroll
    for (int q = 0; q < RD_ROWS; q++) {
      const uint32_t v = word_values(xv[q], is_plink);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const uint32_t g = (v >> (2 * j)) & 0x03030303u;
        acc[q][0] = __dp4a(g, pl[j].x, acc[q][0]);
        acc[q][1] = __dp4a(g, pl[j].y, acc[q][1]);
        acc[q][2] = __dp4a(g, pl[j].z, acc[q][2]);
        acc[q][3] = __dp4a(g, pl[j].w, acc[q][3]);
      }
    }
    if ((it & 65535) == 65535) {  // 32-bit headroom: 65536 x 4 x 3060 < 2^32
#pragma unroll
      for (int q = 0; q < RD_ROWS; q++) {
#pragma unroll
        for (int p = 0; p < 4; p++) { tot[q] += (unsigned long long)acc[q][p] << (8 * p); acc[q][p] = 0; }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < RD_ROWS; q++) {
    unsigned long long a = tot[q];
#pragma unroll
    for (int p = 0; p < 4; p++) a += (unsigned long long)acc[q][p] << (8 * p);
    for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    // integer partials below 2^53: the fp64 atomic sum over the K splits is exact in any order
    if (lane == 0 && r0 + q < rows) atomicAdd(&m[r0 + q], (double)a);
  }
}
// m = X (X^T 1) for a packed matrix x[rows][nb]; d_s: input_row_sums_scratch(nb) unsigned ints
inline size_t input_row_sums_scratch(long nb) { return 2 * (size_t)(4 * nb + 16); }
inline int input_row_sums(const uint8_t *d_x, long rows, long nb, int is_plink, unsigned int *d_s,
                          double *d_m, cudaStream_t stream) {
  int dev_ = 0;
  cudaGetDevice(&dev_);
  const long sms = device_sm_count(dev_);
  MB_CUDA(cudaMemsetAsync(d_s, 0, sizeof(unsigned int) * (4 * nb + 16), stream));
  int ysplit = (int)(rows / 512 < 1 ? 1 : rows / 512 > 64 ? 64 : rows / 512);
  if (nb % 4 == 0 && (reinterpret_cast<uintptr_t>(d_x) & 15) == 0) {
    const long nw = nb / 4;
    ysplit = (int)(ceil_div(rows, 128) > 4096 ? 4096 : ceil_div(rows, 128));  // <= 2 flushes per thread
    snp_colsum_words_kernel<<<dim3((unsigned)ceil_div(nw, 128), ysplit), 128, 0, stream>>>(
        reinterpret_cast<const uint32_t *>(d_x), rows, nw, is_plink, d_s);
    uint4 *planes = reinterpret_cast<uint4 *>(d_s + 4 * nb + 16);  // 16-byte aligned: nb % 4 == 0
    colsum_planes_kernel<<<(unsigned)ceil_div(4 * nw, 256), 256, 0, stream>>>(d_s, nw, planes);
    const long rblocks = ceil_div(rows, 8 * RD_ROWS);
    long ks = ceil_div(sms * 6, rblocks);
    if (ks > ceil_div(nw, 256)) ks = ceil_div(nw, 256);
    if (ks < 1) ks = 1;
    MB_CUDA(cudaMemsetAsync(d_m, 0, sizeof(double) * rows, stream));
    row_dot_words_kernel<<<dim3((unsigned)rblocks, (unsigned)ks), 256, 0, stream>>>(
        reinterpret_cast<const uint32_t *>(d_x), rows, nw, is_plink, planes, d_m);
    count_launch();
  } else {
    snp_colsum_kernel<<<dim3((unsigned)ceil_div(nb, 256), ysplit), 256, 0, stream>>>(d_x, rows, nb,
                                                                                     is_plink, d_s);
    row_dot_kernel<<<(unsigned)ceil_div(rows, 8), 256, 0, stream>>>(d_x, rows, nb, is_plink, d_s, d_m);
  }
  count_launch(2);
  MB_LAUNCH_CHECK();
  return 0;
}
// mode 0 fallback GEMM: dp4a SIMT, 64x64 output tile per block, upper-triangle blocks only.
// C32[i][j] (+)= sum_k X8[i][k] X8[j][k]
__global__ void __launch_bounds__(256)
crossprod_simt_kernel(const int8_t *__restrict__ X8, long kpad, int32_t *__restrict__ C32,
                      long ld32, int accumulate) {
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj < bi) return;
  __shared__ uint32_t As[64][17], Bs[64][17];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 4x4 outputs
  int acc[4][4] = {};
  for (long k0 = 0; k0 < kpad; k0 += 64) {
    __syncthreads();
    for (int e = tid; e < 64 * 16; e += 256) {
      const int r = e >> 4, w = e & 15;
      As[r][w] = *reinterpret_cast<const uint32_t *>(X8 + ((long)bi * 64 + r) * kpad + k0 + w * 4);
      Bs[r][w] = *reinterpret_cast<const uint32_t *>(X8 + ((long)bj * 64 + r) * kpad + k0 + w * 4);
    }
    __syncthreads();
#pragma unroll
    for (int w = 0; w < 16; w++) {
      uint32_t a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; u++) { a[u] = As[ty * 4 + u][w]; b[u] = Bs[tx * 4 + u][w]; }
#pragma unroll
      for (int u = 0; u < 4; u++)
#pragma unroll
        for (int v = 0; v < 4; v++) acc[u][v] = __dp4a((int)a[u], (int)b[v], acc[u][v]);
    }
  }
#pragma unroll
  for (int u = 0; u < 4; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) {
      int32_t *p = C32 + ((long)bi * 64 + ty * 4 + u) * ld32 + (long)bj * 64 + tx * 4 + v;
      *p = accumulate ? *p + acc[u][v] : acc[u][v];
    }
}

}  // namespace mb200
