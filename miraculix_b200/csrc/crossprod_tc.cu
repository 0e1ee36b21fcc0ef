// crossprod_tc.cu -- the crossproduct GEMM  C32 = X8 X8^T  on 5th-gen tensor cores
// (SURVEY.md 8a rows a9/a10).
//
// Replaces the CUTLASS 2.10 uint4b_t tensor-op GEMM with the custom two-pass
// mma.sync.m8n8k32.u4 micro-kernel (reference src/cuda/snp_multiply_cuda.cu:123-171,
// src/cuda/snp_multiply_cuda.h:121-199) and its host-driven tile loop (:190-340).
// tcgen05 has no 4-bit integer MMA, so the operands are int8 (expanded once per K-slab by
// expand_2bit_kernel) and the whole upper triangle is one persistent launch:
//
//   warp 0      TMA producer   cp.async.bulk.tensor 2-D tiles, 128-byte swizzle, mbarrier ring
//   warp 1      MMA issuer     tcgen05.mma.cta_group::1.kind::i8, M=128 N=256 K=32, SS operands,
//                              int32 accumulators in TMEM (2 x 256 columns, double-buffered)
//   warps 2..5  epilogue       tcgen05.ld 32x32b.x32 -> registers -> int32 global (+= for slabs>0)
//
// One CTA per SM, static round-robin over a host-built list of upper-triangle tiles that is
// ordered in L2-friendly super-blocks.
#include <cuda.h>
#include "common.cuh"
#include "handle.h"

namespace mb200 {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 128;  // BK in bytes = int8 elements
constexpr int TC_STAGES = 4;
constexpr int TC_A_BYTES = TC_BM * TC_BK;  // 16 KB
constexpr int TC_B_BYTES = TC_BN * TC_BK;  // 32 KB
constexpr int TC_STAGE_BYTES = TC_A_BYTES + TC_B_BYTES;
constexpr int TC_SMEM = TC_STAGES * TC_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TC_THREADS = 192;

__device__ __forceinline__ double cp_affine(double v, double mr, double mc, double alpha,
                                            double shift, double rc, int do_scale) {
  v = __dadd_rn(v, __dmul_rn(alpha, mr));  // same operation order as grm_affine (crossprod_kernels.cuh)
  v = __dadd_rn(v, __dmul_rn(alpha, mc));
  v = __dadd_rn(v, shift);
  return do_scale ? __dmul_rn(v, rc) : v;
}

struct TileCoord {
  int mi, nj;  // operand tiles: rows mi*128.. of X against rows nj*256.. of X
  int mo;      // output row tile: C32 row = mo*128 + r (== mi unless the output is a compact
               // panel of one partition's row tiles, crossprod.cu)
  // Balanced tail: the tiles of the last, partial round of the static round-robin are cut along K
  // into pieces [kb0, kb1) dealt one per CTA; a piece adds its int32 partial into scratch tile
  // `slot` (red.global.add) and tail_finish_kernel applies the epilogue.  Whole tiles: kb1 = -1.
  int kb0, kb1, slot;
};

__global__ void __launch_bounds__(TC_THREADS, 1)
crossprod_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b,
                    const TileCoord *__restrict__ tiles, int num_tiles, int num_kblocks,
                    int32_t *__restrict__ C32, long ld32, int accumulate, const CpFinish fin) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                              ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + TC_STAGES * TC_STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + TC_STAGES;
  uint64_t *tfull = bars + 2 * TC_STAGES, *tempty = bars + 2 * TC_STAGES + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * TC_STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_STAGES; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; a++) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 0) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer (whole warp walks the protocol, one elected lane issues) =====
    int stage = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const TileCoord tc = tiles[t];
      const int kb1 = tc.kb1 < 0 ? num_kblocks : tc.kb1;
      for (int kb = tc.kb0; kb < kb1; kb++) {
        mbar_wait_warp(&empty[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t *sa = smem + stage * TC_STAGE_BYTES, *sb = sa + TC_A_BYTES;
          mbar_expect_tx(&full[stage], TC_STAGE_BYTES);
          tma_load_2d(sa, &tmap_a, &full[stage], kb * TC_BK, tc.mi * TC_BM);
          tma_load_2d(sb, &tmap_b, &full[stage], kb * TC_BK, tc.nj * TC_BN);
        }
        __syncwarp();
        if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (elect.sync issue: no uniformisation loops around UTCIMMA) =====
    constexpr uint32_t idesc = make_idesc_i8(TC_BM, TC_BN, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      mbar_wait_warp(&tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * TC_BN;
      const int kb0 = tiles[t].kb0, kb1 = tiles[t].kb1 < 0 ? num_kblocks : tiles[t].kb1;
      for (int kb = kb0; kb < kb1; kb++) {
        mbar_wait_warp(&full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * TC_STAGE_BYTES);
          const uint64_t adesc = make_desc_k_sw128(sa);
          const uint64_t bdesc = make_desc_k_sw128(sa + TC_A_BYTES);
#pragma unroll
          for (int k4 = 0; k4 < TC_BK / 32; k4++)  // +32 bytes along K = +2 in the >>4 address field
            mma_i8_ss(d_tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc, (kb > kb0 || k4) ? 1u : 0u);
          tc_commit(&empty[stage]);
          if (kb == kb1 - 1) tc_commit(&tfull[acc]);
        }
        __syncwarp();
        if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quadrant = warp % 4 =====
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    const double alpha = fin.kind ? -1.0 / (double)fin.n : 0.0;
    const double shift = fin.affine ? fin.scal[0] / ((double)fin.n * (double)fin.n) : 0.0;
    const double rc = fin.affine && fin.do_scale ? 1.0 / fin.scal[1] : 1.0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const TileCoord tc = tiles[t];
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const long row = (long)tc.mo * TC_BM + quad * 32 + lane;
      int32_t *crow = C32 + row * ld32 + (long)tc.nj * TC_BN;
      const uint32_t taddr = tmem_base + acc * TC_BN + ((uint32_t)(quad * 32) << 16);
      // global row of the result this thread owns (panels are compact in the row tiles)
      const long gi = fin.kind == 2 ? ((long)tc.mo * fin.nparts + fin.part) * TC_BM + quad * 32 + lane
                                    : row;
      const bool iok = fin.kind && gi < fin.n;
      const double mi = (fin.affine && iok) ? fin.m[gi] : 0.0;
#pragma unroll 1
      for (int c = 0; c < TC_BN / 32; c++) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (tc.slot >= 0) {  // K piece of a tail tile: int32 partial into the scratch tile
          int *dst = fin.tail + ((long)tc.slot * TC_BM + quad * 32 + lane) * TC_BN + c * 32;
#pragma unroll
          for (int q = 0; q < 32; q++) atomicAdd(dst + q, (int)r[q]);
          continue;
        }
        if (fin.kind == 0) {
          int4 *dst = reinterpret_cast<int4 *>(crow + c * 32);
#pragma unroll
          for (int q = 0; q < 8; q++) {
            int4 v = make_int4((int)r[4 * q], (int)r[4 * q + 1], (int)r[4 * q + 2], (int)r[4 * q + 3]);
            if (accumulate) {
              const int4 o = dst[q];
              v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
            }
            dst[q] = v;
          }
          continue;
        }
        if (accumulate) {
          const int4 *src = reinterpret_cast<const int4 *>(crow + c * 32);
#pragma unroll
          for (int q = 0; q < 8; q++) {
            const int4 o = src[q];
            r[4 * q] += (uint32_t)o.x; r[4 * q + 1] += (uint32_t)o.y;
            r[4 * q + 2] += (uint32_t)o.z; r[4 * q + 3] += (uint32_t)o.w;
          }
        }
        const long j0 = (long)tc.nj * TC_BN + c * 32;
        // row sums of this chunk's 32 columns: one per lane (read-only path), broadcast by shuffle
        // -- a plain m[j] load inside the store loop is serialised behind the stores (aliasing)
        const double mlane = (fin.affine && j0 + lane < fin.n) ? __ldg(fin.m + j0 + lane) : 0.0;
        double mjv[32];
#pragma unroll
        for (int q = 0; q < 32; q++) mjv[q] = __shfl_sync(0xffffffffu, mlane, q);
        if (!iok) {
          // padding row: nothing to write (stay converged for the next collective tcgen05.ld)
        } else if (fin.kind == 1) {
          double *col = fin.out + gi + j0 * fin.n;      // (i, j0 + q): lanes = consecutive i
          double *mir = fin.out + j0 + gi * fin.n;      // (j0 + q, i): this thread's own column
#pragma unroll
          for (int q = 0; q < 32; q++) {
            const long j = j0 + q;
            if (j < fin.n && gi <= j) {
              const double v = (double)(int)r[q];
              if (fin.affine) {
                const double mj = mjv[q];
                col[(long)q * fin.n] = cp_affine(v, mi, mj, alpha, shift, rc, fin.do_scale);
                mir[q] = cp_affine(v, mj, mi, alpha, shift, rc, fin.do_scale);
              } else {
                col[(long)q * fin.n] = v;
                mir[q] = v;
              }
            }
          }
        } else {
          double *prow = fin.out + row * fin.n + j0;
#pragma unroll
          for (int q = 0; q < 32; q++) {
            const long j = j0 + q;
            if (j < fin.n) {
              const double v = (double)(int)r[q];
              prow[q] = fin.affine ? cp_affine(v, mi, mjv[q], alpha, shift, rc, fin.do_scale) : v;
            }
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// 2-D u8 tensor [rows][row_bytes] (row pitch = pitch bytes), box = box_rows x box_bytes.
int make_tmap_u8_2d(CUtensorMap *map, const void *base, long row_bytes, long rows, long pitch,
                    int box_bytes, int box_rows, int swizzle128) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return fail(-10, __func__, __LINE__, "cuTensorMapEncodeTiled not available");
  cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch};
  cuuint32_t box[2] = {(cuuint32_t)box_bytes, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(-11, __func__, __LINE__, "cuTensorMapEncodeTiled -> %d", (int)r);
  return 0;
}

// Upper-triangle tile list in super-blocks of 16 x 8 tiles (2048 x 2048 outputs) so that the
// ~148 tiles in flight share operand panels in L2.
// part/nparts: row tile mi belongs to partition mi % nparts and is written to the compact output
// row tile mi / nparts (2-D output-tile partition over GPUs: no communication, cyclic = balanced).
// Epilogue of the K-split tail tiles: scratch int32 tile -> the same finish as the GEMM epilogue.
__global__ void __launch_bounds__(256)
tail_finish_kernel(const TileCoord *__restrict__ tail_tiles, const int *__restrict__ scratch,
                   const CpFinish fin) {
  const TileCoord tc = tail_tiles[blockIdx.x];
  const double alpha = -1.0 / (double)fin.n;
  const double shift = fin.affine ? fin.scal[0] / ((double)fin.n * (double)fin.n) : 0.0;
  const double rc = fin.affine && fin.do_scale ? 1.0 / fin.scal[1] : 1.0;
  const int *src = scratch + (long)blockIdx.x * TC_BM * TC_BN;
  for (int e = threadIdx.x; e < TC_BM * TC_BN; e += 256) {
    // kind 1 stores column-major (consecutive threads = consecutive rows), kind 2 row-major
    const int r = fin.kind == 1 ? e % TC_BM : e / TC_BN, c = fin.kind == 1 ? e / TC_BM : e % TC_BN;
    const long lrow = (long)tc.mo * TC_BM + r;
    const long gi = fin.kind == 2 ? ((long)tc.mo * fin.nparts + fin.part) * TC_BM + r : lrow;
    const long j = (long)tc.nj * TC_BN + c;
    if (gi >= fin.n || j >= fin.n) continue;
    const double v = (double)src[r * TC_BN + c];
    const double mi = fin.affine ? fin.m[gi] : 0.0, mj = fin.affine ? fin.m[j] : 0.0;
    if (fin.kind == 1) {
      if (gi > j) continue;
      fin.out[gi + j * fin.n] = fin.affine ? cp_affine(v, mi, mj, alpha, shift, rc, fin.do_scale) : v;
      fin.out[j + gi * fin.n] = fin.affine ? cp_affine(v, mj, mi, alpha, shift, rc, fin.do_scale) : v;
    } else {
      fin.out[lrow * fin.n + j] = fin.affine ? cp_affine(v, mi, mj, alpha, shift, rc, fin.do_scale) : v;
    }
  }
}

// split_kblocks > 0: K blocks per tile of a single-slab product whose last partial round may be
// cut into K pieces (see TileCoord); *num_tail = tiles cut, their coordinates follow the work list
// at (*d_tiles)[*num_tiles ...].
int build_tile_list(long rows, int part, int nparts, TileCoord **d_tiles, int *num_tiles,
                    cudaStream_t stream, int split_kblocks, int grid_max, int *num_tail,
                    int *real_tiles) {
  const int Mt = (int)ceil_div(rows, TC_BM), Nt = (int)ceil_div(rows, TC_BN);
  std::vector<TileCoord> v;
  for (int sn = 0; sn < Nt; sn += 8)
    for (int sm = 0; sm < Mt; sm += 16 * nparts)
      for (int nj = sn; nj < sn + 8 && nj < Nt; nj++)
        for (int mi = sm + part; mi < sm + 16 * nparts && mi < Mt; mi += nparts)
          if ((long)TC_BN * (nj + 1) > (long)TC_BM * mi) v.push_back({mi, nj, mi / nparts, 0, -1, -1});
  int ntail = 0;
  const int T = (int)v.size();
  if (split_kblocks > 0 && grid_max > 0 && T > grid_max) {
    const int rem = T % grid_max;
    const int P = rem ? grid_max / rem : 0;              // K pieces per tail tile
    if (rem && P >= 2 && split_kblocks >= 4 * P) {       // worth it: <= half a round, pieces >= 4 blocks
      std::vector<TileCoord> tail(v.end() - rem, v.end());
      v.resize(T - rem);
      for (int p = 0; p < P; p++)                        // piece-major: the pieces of one tile
        for (int t = 0; t < rem; t++) {                  // run on different CTAs at the same time
          TileCoord c = tail[t];
          c.kb0 = (int)((long)split_kblocks * p / P);
          c.kb1 = (int)((long)split_kblocks * (p + 1) / P);
          c.slot = t;
          v.push_back(c);
        }
      ntail = rem;
      *num_tiles = (int)v.size();
      v.insert(v.end(), tail.begin(), tail.end());
    }
  }
  if (!ntail) *num_tiles = T;
  if (num_tail) *num_tail = ntail;
  if (real_tiles) *real_tiles = T;
  if (v.empty()) v.push_back({0, 0, 0, 0, -1, -1});  // keep the allocation non-empty; the launch is skipped
  MB_CUDA(cudaMallocAsync(d_tiles, sizeof(TileCoord) * v.size(), stream));
  MB_CUDA(cudaMemcpyAsync(*d_tiles, v.data(), sizeof(TileCoord) * v.size(), cudaMemcpyHostToDevice,
                          stream));
  MB_CUDA(cudaStreamSynchronize(stream));  // v goes out of scope
  return 0;
}

int crossprod_tc_launch(const int8_t *X8, long rows_pad, long kcols, long pitch,
                        const TileCoord *d_tiles,
                        int num_tiles, int32_t *C32, long ld32, int accumulate, int sm_count,
                        cudaStream_t stream, const CpFinish *finish, int num_tail) {
  MB_CUDA(cudaFuncSetAttribute(crossprod_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               TC_SMEM));
  CUtensorMap ta, tb;
  MB_CHECK(make_tmap_u8_2d(&ta, X8, kcols, rows_pad, pitch, TC_BK, TC_BM, 1));
  MB_CHECK(make_tmap_u8_2d(&tb, X8, kcols, rows_pad, pitch, TC_BK, TC_BN, 1));
  if (num_tiles <= 0) return 0;
  const int grid = num_tiles < sm_count ? num_tiles : sm_count;
  CpFinish fin = finish ? *finish : CpFinish();
  int *scratch = nullptr;
  if (num_tail > 0) {
    if (!finish || !finish->kind) return fail(-2, __func__, __LINE__, "split tail needs a finishing epilogue");
    const size_t bytes = sizeof(int) * (size_t)num_tail * TC_BM * TC_BN;
    MB_CUDA(cudaMallocAsync(&scratch, bytes, stream));
    MB_CUDA(cudaMemsetAsync(scratch, 0, bytes, stream));
    fin.tail = scratch;
  }
  crossprod_tc_kernel<<<grid, TC_THREADS, TC_SMEM, stream>>>(ta, tb, d_tiles, num_tiles,
                                                             (int)(kcols / TC_BK), C32, ld32,
                                                             accumulate, fin);
  count_launch();
  MB_LAUNCH_CHECK();
  if (num_tail > 0) {
    tail_finish_kernel<<<num_tail, 256, 0, stream>>>(d_tiles + num_tiles, scratch, fin);
    count_launch();
    MB_LAUNCH_CHECK();
    MB_CUDA(cudaFreeAsync(scratch, stream));
  }
  return 0;
}

}  // namespace mb200
