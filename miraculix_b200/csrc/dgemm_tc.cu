// dgemm_tc.cu -- dgemm_compressed on the 5th-gen tensor cores, exact int8 digit slices (mode 1).
//
// Why: at ncol = 10 the fp64 FMA formulation needs 2*snps*indiv*ncol fp64 flops = 54 ms of B200
// fp64 pipe for configs[2], 14x the 3.8 ms the packed matrix needs to stream from HBM once.  The
// genotypes are exact small integers, so each fp64 column of B is split into S signed 8-bit
// digits of a per-column fixed-point number (|q| <= 2^(8S-2)); the digit products are EXACT in
// int32 on tcgen05 kind::i8 and are recombined exactly in the epilogue:
//     sum_k z_rk b_kc  =  2^(e_c - P) * sum_j 256^j * ( sum_k z_rk d_kcj )          P = 8S - 2
// The only error is the quantisation of B, |db| <= 2^(e_c - P - 1) <= 2^-54 max|b_c| at S = 7
// (fp64 itself carries 2^-53 relative per element).
//
// Replaces the reference's CUTLASS SIMT u2 x f64 GEMM (src/cuda/dgemm_compressed_cuda.cu:328-385,
// micro-kernel dgemm_compressed_cuda.h:253-268).  Kernels:
//   colmax_kernel     amax_c = max_k |B[k,c]|
//   quantize_kernel   Bq[(c*S + j)][k] = digit j of rint(B[k,c] * 2^(P - e_c))       (int8, K-major)
//   dgemm_tc_kernel   persistent, warp-specialised, one CTA per SM:
//       warps 16,18  TMA producers: packed Z tiles [128 rows x 128 B] (swizzle 128B); Bq tiles
//       warps 0..15  expanders: 2-bit -> int8 (7 integer ops per 16 genotypes thanks to the
//                    interleaved layout), tcgen05.st into the TMEM A-operand ring; epilogue
//       warp 17      MMA issuer: tcgen05.mma kind::i8, M=128, N=16..80, K=32, A from TMEM,
//                    B (digits) from shared memory, int32 accumulators in TMEM
//   HBM-bound by design: algorithmic bytes = the packed matrix once (+ B, C).
#include <cuda.h>
#include <math.h>
#include <stdlib.h>
#include "common.cuh"
#include "handle.h"

namespace mb200 {

int make_tmap_u8_2d(CUtensorMap *map, const void *base, long row_bytes, long rows, long pitch,
                    int box_bytes, int box_rows, int swizzle128);
int dgemm_center_init(Handle *h, int trans, int n, const double *dB, long ldb, double *dC,
                      long ldc, int centered);

bool dgemm_tc_available() { return true; }

// ---------------------------------------------------------------------------------------------
// pre-pass: per-column amax and digit slicing
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colmax_kernel(const double *__restrict__ B, long ldb, long k, unsigned long long *__restrict__ amax) {
  const int col = blockIdx.x;
  const double *b = B + (size_t)col * ldb;
  double m = 0.0;
  for (long j = (long)blockIdx.y * 256 + threadIdx.x; j < k; j += (long)gridDim.y * 256)
    m = fmax(m, fabs(b[j]));
  for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  // non-negative doubles order like their bit patterns
  if ((threadIdx.x & 31) == 0) atomicMax(&amax[col], (unsigned long long)__double_as_longlong(m));
}

// One thread per k; writes S*ncols rows of Bq (coalesced along k).  Rows >= ncols*S and the
// K padding stay as initialised (zero).
__global__ void __launch_bounds__(256)
quantize_kernel(const double *__restrict__ B, long ldb, long k, int ncols, int S,
                const unsigned long long *__restrict__ amax, int8_t *__restrict__ Bq, long kpitch,
                int *__restrict__ exps) {
  const long j = (long)blockIdx.x * 256 + threadIdx.x;
  const int P = 8 * S - 2;
  for (int c = 0; c < ncols; c++) {
    const double am = __longlong_as_double((long long)amax[c]);
    int e = 0;
    if (am > 0.0) frexp(am, &e);  // am = m 2^e, m in [0.5, 1)  ->  am < 2^e
    if (j == 0) exps[c] = e;
    if (j < k) {
      const double b = B[j + (size_t)c * ldb];
      long long q = am > 0.0 ? llrint(scalbn(b, P - e)) : 0;  // |q| <= 2^P
      int8_t *dst = Bq + (size_t)(c * S) * kpitch + j;
      for (int s = 0; s < S - 1; s++) {
        const int d = (int)((q + 128) & 255) - 128;  // balanced digit in [-128, 127]
        dst[(size_t)s * kpitch] = (int8_t)d;
        q = (q - d) >> 8;
      }
      dst[(size_t)(S - 1) * kpitch] = (int8_t)q;  // |top digit| <= 2^(P - 8(S-1)) + 1 = 65
    }
  }
}

// ---------------------------------------------------------------------------------------------
// main kernel
// ---------------------------------------------------------------------------------------------
constexpr int DT_RT = 4;                   // row tiles (of 128 rows) per CTA = TMEM accumulators
constexpr int DT_NMAX = 80;                // max MMA N (digit rows) per pass
constexpr int DT_ZSLOTS = 8;               // packed-Z ring: [128 rows][128 B] = 16 KB each
constexpr int DT_ZBYTES = 128 * 128;
constexpr int DT_BSLOTS = 4;               // digit ring: 2 sub-tiles [N rows][128 B] per slot (K = 256)
constexpr int DT_ASLOTS = 6;               // TMEM A-operand ring: 32 columns (K = 128) each
constexpr int DT_ACOL0 = DT_RT * DT_NMAX;  // 320: accumulators below, A ring above
constexpr int DT_THREADS = 19 * 32;
constexpr int DT_NBAR = 2 * DT_ZSLOTS + 2 * DT_BSLOTS + 2 * DT_ASLOTS + 2;

__host__ __device__ constexpr int dt_smem_bytes(int N) {
  return DT_ZSLOTS * DT_ZBYTES + DT_BSLOTS * 2 * N * 128 + 1024 /*align*/ + 8 * DT_NBAR + 16;
}

struct DtParams {
  long rows;            // rows of the packed matrix used (indiv for 'N', snps for 'T')
  int nchunks;          // K chunks of 512 genotypes (= 128 packed bytes per row)
  int row_blocks;       // ceil(rows / 512)
  int ksplits, cps;     // K splits and chunks per split
  int N, S, ncols;      // MMA N (multiple of 16), digits per column, columns this pass
  double *C;            // centre-initialised output, column-major
  long ldc;
  const int *exps;      // e_c per column
  int dbg;              // TRACE builds only: 1 skip tcgen05.st, 2 skip MMAs, 4 skip LDS of Z, 8 skip ALU
};

// Optional cycle accounting (MB200_TRACE=1): CTA 0 accumulates clock64() segments of expander
// warp 0 and of the MMA warp into g_dt_trace; read back with mb200_debug_read_trace.
__device__ long long g_dt_trace[32];
// timeline of ring slots [DT_EV0, DT_EV0 + 256) of CTA 0: [0] expander passes a_empty, [1] expander
// arrives a_full, [2] MMA warp passes a_full, [3] MMA warp done issuing, [4] z_full passed (per chunk)
constexpr unsigned DT_EV0 = 4096;
__device__ long long g_dt_events[5][256];
#define DT_SEG(slot)                                   \
  if (TRACE) {                                         \
    const long long _now = clock64();                  \
    tr[slot] += _now - tlast;                          \
    tlast = _now;                                      \
  }

template <bool TRACE>
__global__ void __launch_bounds__(DT_THREADS, 1)
dgemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_z,
                const __grid_constant__ CUtensorMap tmap_b, const DtParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                              ~(uintptr_t)1023);
  uint8_t *zring = smem;
  uint8_t *bring = smem + DT_ZSLOTS * DT_ZBYTES;
  const int sub_bytes = p.N * 128;          // one digit sub-tile: N rows x 128 K-bytes
  const int bslot_bytes = 2 * sub_bytes;    // half a K chunk: 256
  uint64_t *bars = reinterpret_cast<uint64_t *>(bring + DT_BSLOTS * bslot_bytes);
  uint64_t *z_full = bars, *z_empty = z_full + DT_ZSLOTS;
  uint64_t *b_full = z_empty + DT_ZSLOTS, *b_empty = b_full + DT_BSLOTS;
  uint64_t *a_full = b_empty + DT_BSLOTS, *a_empty = a_full + DT_ASLOTS;
  uint64_t *d_full = a_empty + DT_ASLOTS, *d_empty = d_full + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(d_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_units = p.row_blocks * p.ksplits;

  if (threadIdx.x == 0) {
    for (int i = 0; i < DT_ZSLOTS; i++) { mbar_init(&z_full[i], 1); mbar_init(&z_empty[i], 4); }
    for (int i = 0; i < DT_BSLOTS; i++) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < DT_ASLOTS; i++) { mbar_init(&a_full[i], 4); mbar_init(&a_empty[i], 1); }
    mbar_init(d_full, 1);
    mbar_init(d_empty, 16);
    fence_mbar_init();
    tma_prefetch_desc(&tmap_z);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 17) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 16) {
    // ===== TMA producer of the packed matrix (whole warp walks, one elected lane issues) =====
    int zs = 0;
    uint32_t zph = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int rb = u % p.row_blocks, ks = u / p.row_blocks;
      const int c0 = ks * p.cps, c1 = min(c0 + p.cps, p.nchunks);
      for (int kc = c0; kc < c1; kc++) {
#pragma unroll
        for (int g = 0; g < DT_RT; g++) {
          mbar_wait_warp(&z_empty[zs], zph ^ 1);
          if (elect_one()) {
            mbar_expect_tx(&z_full[zs], DT_ZBYTES);
            tma_load_2d(zring + zs * DT_ZBYTES, &tmap_z, &z_full[zs], kc * 128,
                        rb * (DT_RT * 128) + g * 128);
          }
          __syncwarp();
          if (++zs == DT_ZSLOTS) { zs = 0; zph ^= 1; }
        }
      }
    }
  } else if (warp == 18) {
    // ===== TMA producer of the digit panel (its own warp: the two rings never block each other)
    int bs = 0;
    uint32_t bph = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int ks = u / p.row_blocks;
      const int c0 = ks * p.cps, c1 = min(c0 + p.cps, p.nchunks);
      for (int kh = 2 * c0; kh < 2 * c1; kh++) {  // half chunks of K = 256
        mbar_wait_warp(&b_empty[bs], bph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&b_full[bs], (uint32_t)bslot_bytes);
          tma_load_2d(bring + bs * bslot_bytes, &tmap_b, &b_full[bs], kh * 256, 0);
          tma_load_2d(bring + bs * bslot_bytes + sub_bytes, &tmap_b, &b_full[bs], kh * 256 + 128, 0);
        }
        __syncwarp();
        if (++bs == DT_BSLOTS) { bs = 0; bph ^= 1; }
      }
    }
  } else if (warp == 17) {
    // ===== MMA issuer: the whole warp walks the protocol, one elected lane issues (a lane==0
    // branch makes ptxas wrap every UTCIMMA in a uniformisation loop: 60 instead of 45 cycles
    // per MMA, scripts/probes/mma_probe.cu) =====
    const uint32_t idesc = make_idesc_i8(128, p.N, 1, 1);
    int as = 0, bs = 0;
    uint32_t aph = 0, bph = 0, dph = 0;
    long long tr[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = TRACE ? clock64() : 0;
    bool a_ready = false, b_ready = false;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int ks = u / p.row_blocks;
      const int c0 = ks * p.cps, c1 = min(c0 + p.cps, p.nchunks);
      mbar_wait_warp(d_empty, dph ^ 1);  // epilogue of the previous unit has drained TMEM
      tc_fence_after();
      DT_SEG(0)
      for (int kc = c0; kc < c1; kc++) {
#pragma unroll
        for (int sub = 0; sub < 4; sub++) {
          if ((sub & 1) == 0) {
            if (!b_ready) mbar_wait_warp(&b_full[bs], bph);
            tc_fence_after();
            DT_SEG(1)
          }
          const uint64_t bdesc =
              make_desc_k_sw128(smem_u32(bring + bs * bslot_bytes) + (sub & 1) * sub_bytes);
#pragma unroll
          for (int g = 0; g < DT_RT; g++) {
            if (!a_ready) mbar_wait_warp(&a_full[as], aph);
            tc_fence_after();
            DT_SEG(2)
            // A try_wait costs ~90 cycles even on a completed barrier (scripts/probes/
            // instr_probe.cu): poll the NEXT ring slots now, consume the answer after the MMAs
            // of this slot have been issued, so the latency hides behind their execution.
            const int as_n = as + 1 == DT_ASLOTS ? 0 : as + 1;
            const uint32_t aph_n = as + 1 == DT_ASLOTS ? aph ^ 1 : aph;
            const bool a_poll = mbar_try_wait(&a_full[as_n], aph_n);
            const bool last_of_b = (sub & 1) && g == DT_RT - 1;
            const int bs_n = bs + 1 == DT_BSLOTS ? 0 : bs + 1;
            const uint32_t bph_n = bs + 1 == DT_BSLOTS ? bph ^ 1 : bph;
            bool b_poll = false;
            if (last_of_b) b_poll = mbar_try_wait(&b_full[bs_n], bph_n);
            DT_SEG(4)
            if (elect_one()) {
              const uint32_t d_tmem = tmem_base + g * p.N;
              const uint32_t a_tmem = tmem_base + DT_ACOL0 + as * 32;
#pragma unroll
              for (int k4 = 0; k4 < 4; k4++)  // K = 32 per MMA: 8 TMEM columns of A, 32 B of Bq
                if (!TRACE || !(p.dbg & 2))
                mma_i8_ts(d_tmem, a_tmem + 8 * k4, bdesc + 2 * k4, idesc,
                          (kc > c0 || sub > 0 || k4 > 0) ? 1u : 0u);
              DT_SEG(5)
              tc_commit(&a_empty[as]);
              if (last_of_b) tc_commit(&b_empty[bs]);
              DT_SEG(6)
            }
            __syncwarp();
            DT_SEG(7)
            a_ready = __all_sync(0xffffffffu, a_poll);
            if (last_of_b) b_ready = __all_sync(0xffffffffu, b_poll);
            DT_SEG(3)
            as = as_n;
            aph = aph_n;
          }
          if (sub & 1) {
            bs = bs + 1 == DT_BSLOTS ? 0 : bs + 1;
            if (bs == 0) bph ^= 1;
          }
        }
      }
      if (elect_one()) tc_commit(d_full);
      __syncwarp();
      dph ^= 1;
    }
    if (TRACE && blockIdx.x == 0 && lane == 0)
      for (int i = 0; i < 8; i++) g_dt_trace[16 + i] = tr[i];
  } else {
    // ===== expanders (and epilogue): group g owns row tile g, thread = row = TMEM lane =====
    const int g = warp >> 2, w = warp & 3;
    const int r = 32 * w + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(32 * w) << 16);
    const int P = 8 * p.S - 2;
    uint32_t cnt = 0;  // K chunks processed by this CTA so far
    uint32_t dph = 0;
    long long tr[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = TRACE ? clock64() : 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int rb = u % p.row_blocks, ks = u / p.row_blocks;
      const int c0 = ks * p.cps, c1 = min(c0 + p.cps, p.nchunks);
      for (int kc = c0; kc < c1; kc++, cnt++) {
        const uint32_t zseq = cnt * DT_RT + g;
        const int zs = zseq % DT_ZSLOTS;
        mbar_wait(&z_full[zs], (zseq / DT_ZSLOTS) & 1);
        DT_SEG(0)
        if (TRACE && blockIdx.x == 0 && w == 0 && lane == 0 && cnt * 16 + g - DT_EV0 < 256u)
          g_dt_events[4][cnt * 16 + g - DT_EV0] = clock64();
        const uint8_t *zrow = zring + zs * DT_ZBYTES + r * 128;
        uint4 v[8];
#pragma unroll
        for (int c = 0; c < 8; c++)  // 128-byte swizzle: 16-byte chunk index ^ (row & 7)
          if (!TRACE || !(p.dbg & 4))
            v[c] = *reinterpret_cast<const uint4 *>(zrow + ((c ^ (r & 7)) << 4));
          else
            v[c] = make_uint4(cnt, r, c, kc);
        __syncwarp();
        if (lane == 0) mbar_arrive(&z_empty[zs]);
        DT_SEG(1)
#pragma unroll
        for (int sub = 0; sub < 4; sub++) {
          const uint32_t t = (cnt * 4 + sub) * DT_RT + g;
          const int as = t % DT_ASLOTS;
          mbar_wait(&a_empty[as], ((t / DT_ASLOTS) & 1) ^ 1);
          tc_fence_after();
          DT_SEG(2)
          if (TRACE && blockIdx.x == 0 && w == 0 && lane == 0 && t - DT_EV0 < 256u)
            g_dt_events[0][t - DT_EV0] = clock64();
          const uint32_t wds[8] = {v[2 * sub].x, v[2 * sub].y, v[2 * sub].z, v[2 * sub].w,
                                   v[2 * sub + 1].x, v[2 * sub + 1].y, v[2 * sub + 1].z,
                                   v[2 * sub + 1].w};
          uint32_t a[32];
#pragma unroll
          for (int q = 0; q < 8; q++)
#pragma unroll
            for (int j = 0; j < 4; j++)
              a[4 * q + j] = (TRACE && (p.dbg & 8)) ? wds[q] : (wds[q] >> (2 * j)) & 0x03030303u;
          if (!TRACE || !(p.dbg & 1))
          tmem_st_32x32b_x32(lane_base + DT_ACOL0 + as * 32, a);
          DT_SEG(3)
          tmem_st_wait();
          DT_SEG(4)
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&a_full[as]);
          if (TRACE && blockIdx.x == 0 && w == 0 && lane == 0 && t - DT_EV0 < 256u)
            g_dt_events[1][t - DT_EV0] = clock64();
          DT_SEG(5)
        }
      }
      // ---- epilogue of the unit: exact recombination of the digit sums, fp64 atomics into C
      mbar_wait(d_full, dph);
      dph ^= 1;
      tc_fence_after();
      DT_SEG(6)
      const long row = (long)rb * (DT_RT * 128) + g * 128 + r;
      for (int c = 0; c < p.ncols; c++) {
        uint32_t d[8];
        tmem_ld_32x32b_x8(lane_base + g * p.N + c * p.S, d);
        tmem_ld_wait();
        long long lo = 0, hi = 0;  // sum_j 256^j d_j = hi 2^32 + lo, exact in 64-bit integers
#pragma unroll
        for (int j = 0; j < 4; j++)
          if (j < p.S) lo += (long long)(int)d[j] << (8 * j);
#pragma unroll
        for (int j = 4; j < 8; j++)
          if (j < p.S) hi += (long long)(int)d[j] << (8 * (j - 4));
        const double val = fma((double)hi, 4294967296.0, (double)lo);
        if (row < p.rows) atomicAdd(&p.C[row + (size_t)c * p.ldc], scalbn(val, p.exps[c] - P));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(d_empty);
      DT_SEG(7)
    }
    if (TRACE && blockIdx.x == 0 && threadIdx.x == 0)
      for (int i = 0; i < 8; i++) g_dt_trace[i] = tr[i];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 17) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int ensure_bq(Handle *h) {
  const long kp = (h->pitch_s > h->pitch_i ? h->pitch_s : h->pitch_i) * 4;  // K pitch in int8
  const size_t need = (size_t)DT_NMAX * kp;
  if (h->d_Bq && h->bq_bytes >= need) return 0;
  if (h->d_Bq) MB_CUDA(cudaFree(h->d_Bq));
  h->d_Bq = nullptr;
  MB_CUDA(cudaMalloc(&h->d_Bq, need));
  MB_CUDA(cudaMemsetAsync(h->d_Bq, 0, need, h->stream));
  h->bq_bytes = need;
  return 0;
}

int dgemm_tc_device(Handle *h, int trans, int n, const double *dB, long ldb, double *dC, long ldc) {
  const Options &o = options();
  int S = o.dgemm_slices;
  if (S < 3) S = 3;
  if (S > 8) S = 8;
  const uint8_t *Z = trans ? h->d_snp_major : h->d_indiv_major;
  const long pitch = trans ? h->pitch_s : h->pitch_i;
  const long rows = trans ? h->snps : h->indiv, K = trans ? h->indiv : h->snps;
  const long kpitch = pitch * 4;
  MB_CHECK(ensure_bq(h));
  MB_CHECK(dgemm_center_init(h, trans, n, dB, ldb, dC, ldc, o.centered));

  // scratch inside d_w (4 * (max_n + 16) doubles): [0, max_n) centring, then amax bits, exponents
  unsigned long long *d_amax = reinterpret_cast<unsigned long long *>(h->d_w + (h->max_n + 16));
  int *d_exps = reinterpret_cast<int *>(h->d_w + 2 * (h->max_n + 16));

  static bool attr_set = false;
  static bool trace = false;
  if (!attr_set) {
    MB_CUDA(cudaFuncSetAttribute(dgemm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 dt_smem_bytes(DT_NMAX)));
    MB_CUDA(cudaFuncSetAttribute(dgemm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 dt_smem_bytes(DT_NMAX)));
    const char *e = getenv("MB200_TRACE");
    trace = e && atoi(e) > 0;
    attr_set = true;
  }
  CUtensorMap tz;
  MB_CHECK(make_tmap_u8_2d(&tz, Z, pitch, rows, pitch, 128, 128, 1));

  const int cpp = DT_NMAX / S;  // columns per pass
  const int ev = (int)(h->ev_count % Handle::kEvRing);
  h->ev_trans[ev] = trans;
  bool first = true;
  for (int c0 = 0; c0 < n; c0 += cpp) {
    const int nc = n - c0 < cpp ? n - c0 : cpp;
    const int N = (int)round_up((long)nc * S, 16);
    const double *B = dB + (size_t)c0 * ldb;
    MB_CUDA(cudaMemsetAsync(d_amax, 0, sizeof(unsigned long long) * nc, h->stream));
    int ysplit = (int)ceil_div(K, 256 * 32);
    if (ysplit > 128) ysplit = 128;
    if (ysplit < 1) ysplit = 1;
    colmax_kernel<<<dim3(nc, ysplit), 256, 0, h->stream>>>(B, ldb, K, d_amax);
    quantize_kernel<<<(unsigned)ceil_div(K, 256), 256, 0, h->stream>>>(B, ldb, K, nc, S, d_amax,
                                                                        h->d_Bq, kpitch, d_exps);
    count_launch(2);
    MB_LAUNCH_CHECK();

    CUtensorMap tb;
    MB_CHECK(make_tmap_u8_2d(&tb, h->d_Bq, kpitch, N, kpitch, 128, N, 1));
    DtParams p;
    p.rows = rows;
    p.nchunks = (int)(pitch / 128);
    p.row_blocks = (int)ceil_div(rows, DT_RT * 128);
    long ksplits = ceil_div(20L * h->sm_count, p.row_blocks);
    if (ksplits > p.nchunks / 8) ksplits = p.nchunks / 8;
    if (ksplits < 1) ksplits = 1;
    p.cps = (int)ceil_div(p.nchunks, ksplits);
    p.ksplits = (int)ceil_div(p.nchunks, p.cps);
    p.N = N;
    p.S = S;
    p.ncols = nc;
    p.C = dC + (size_t)c0 * ldc;
    p.ldc = ldc;
    p.exps = d_exps;
    p.dbg = 0;
    if (trace) { const char *e = getenv("MB200_DBG"); p.dbg = e ? atoi(e) : 0; }
    const long units = (long)p.row_blocks * p.ksplits;
    const int grid = (int)(units < h->sm_count ? units : h->sm_count);
    if (first) MB_CUDA(cudaEventRecord(h->ev_begin[ev], h->stream));
    first = false;
    if (trace) dgemm_tc_kernel<true><<<grid, DT_THREADS, dt_smem_bytes(N), h->stream>>>(tz, tb, p);
    else dgemm_tc_kernel<false><<<grid, DT_THREADS, dt_smem_bytes(N), h->stream>>>(tz, tb, p);
    count_launch();
    MB_LAUNCH_CHECK();
  }
  MB_CUDA(cudaEventRecord(h->ev_end[ev], h->stream));
  h->ev_count++;
  return 0;
}

int dgemm_tc_read_trace(long long *out, int n) {
  if (n <= 32) {
    MB_CUDA(cudaMemcpyFromSymbol(out, g_dt_trace, sizeof(long long) * n));
  } else {  // 32 counters followed by the 5 x 256 event timeline
    MB_CUDA(cudaMemcpyFromSymbol(out, g_dt_trace, sizeof(long long) * 32));
    const int m = n - 32 < 5 * 256 ? n - 32 : 5 * 256;
    MB_CUDA(cudaMemcpyFromSymbol(out + 32, g_dt_events, sizeof(long long) * m));
  }
  return 0;
}

}  // namespace mb200
