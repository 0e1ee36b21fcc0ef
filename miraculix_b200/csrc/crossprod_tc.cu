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
#include <mutex>
#include "common.cuh"
#include "handle.h"

namespace mb200 {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 128;  // BK in bytes = int8 elements
constexpr int TC_A_BYTES = TC_BM * TC_BK;  // 16 KB
constexpr int TC_B_BYTES = TC_BN * TC_BK;  // 32 KB
// One CTA per tile: 4 stages of A + B = 48 KB.  CTA pair (cta_group::2, a 256 x 256 tile over two
// SMs): every CTA stages its own 128 rows of A and HALF of B (128 of the 256 rows) = 32 KB, 6
// stages -- per SM a third less L2 -> smem traffic and operand reads per MMA.
template <bool PAIR> struct TcGeom {
  static constexpr int STAGES = PAIR ? 6 : 4;
  static constexpr int STAGE_BYTES = TC_A_BYTES + (PAIR ? TC_B_BYTES / 2 : TC_B_BYTES);
};
constexpr int TC_STG = 32 * 33;  // doubles of one epilogue warp's 32 x 32 staging tile (padded)
constexpr int TC_SMEM = 4 * (TC_A_BYTES + TC_B_BYTES) + 1024 /*align*/ + 256 /*barriers*/ +
                        4 * TC_STG * 8 /*epilogue staging*/;
static_assert(TC_SMEM <= 232448, "shared memory budget");
constexpr int TC_THREADS = 192;

struct TileCoord {
  int mi, nj;  // operand tiles: rows mi*128.. of X against rows nj*256.. of X
  int mo;      // output row tile: C32 row = mo*128 + r (== mi unless the output is a compact
               // panel of one partition's row tiles, crossprod.cu)
  // Balanced tail: the tiles of the last, partial round of the static round-robin are cut along K
  // into pieces [kb0, kb1) dealt one per CTA; a piece adds its int32 partial into scratch tile
  // `slot` (red.global.add) and tail_finish_kernel applies the epilogue.  Whole tiles: kb1 = -1.
  int kb0, kb1, slot;
};

template <bool PAIR>
__global__ void __launch_bounds__(TC_THREADS, 1)
crossprod_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b,
                    const TileCoord *__restrict__ tiles, int num_tiles, int num_kblocks,
                    int32_t *__restrict__ C32, long ld32, int accumulate, const CpFinish fin,
                    int mstride) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                              ~(uintptr_t)1023);
  constexpr int TC_STAGES = TcGeom<PAIR>::STAGES, TC_STAGE_BYTES = TcGeom<PAIR>::STAGE_BYTES;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + TC_STAGES * TC_STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + TC_STAGES;
  uint64_t *tfull = bars + 2 * TC_STAGES, *tempty = bars + 2 * TC_STAGES + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * TC_STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // PAIR: the two CTAs of a cluster share one work item (a pair of row tiles mi, mi + mstride
  // against column tile nj; mstride = 1, or the partition count when the row tiles of one GPU's
  // panel are dealt cyclically); rank 0 (the leader) issues the MMAs for both, every CTA loads its
  // own operands and drains its own 128 accumulator rows.
  const int rank = PAIR ? (int)cluster_ctarank() : 0;
  const int unit0 = PAIR ? blockIdx.x >> 1 : blockIdx.x, nunits = PAIR ? gridDim.x >> 1 : gridDim.x;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_STAGES; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; a++) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], PAIR ? 8 : 4);  // one arrive per epilogue warp (of both CTAs)
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 0) {
    if (PAIR) tmem_alloc_pair(tmem_slot, 512);
    else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers exist before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer (whole warp walks the protocol, one elected lane issues) =====
    int stage = 0;
    uint32_t phase = 0;
    for (int t = unit0; t < num_tiles; t += nunits) {
      const TileCoord tc = tiles[t];
      const int kb1 = tc.kb1 < 0 ? num_kblocks : tc.kb1;
      for (int kb = tc.kb0; kb < kb1; kb++) {
        mbar_wait_warp(&empty[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t *sa = smem + stage * TC_STAGE_BYTES, *sb = sa + TC_A_BYTES;
          if (!PAIR) {
            mbar_expect_tx(&full[stage], TC_STAGE_BYTES);
            tma_load_2d(sa, &tmap_a, &full[stage], kb * TC_BK, tc.mi * TC_BM);
            tma_load_2d(sb, &tmap_b, &full[stage], kb * TC_BK, tc.nj * TC_BN);
          } else {
            // all bytes of both CTAs complete on the LEADER's full barrier; tmap_a's 128-row box
            // serves the A tile and this CTA's half of the B tile
            if (rank == 0) mbar_expect_tx(&full[stage], 2 * TC_STAGE_BYTES);
            const uint32_t lead_full = mapa_u32(smem_u32(&full[stage]), 0);
            tma_load_2d_pair(sa, &tmap_a, lead_full, kb * TC_BK, (tc.mi + rank * mstride) * TC_BM);
            tma_load_2d_pair(sb, &tmap_a, lead_full, kb * TC_BK, tc.nj * TC_BN + rank * (TC_BN / 2));
          }
        }
        __syncwarp();
        if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (elect.sync issue: no uniformisation loops around UTCIMMA) =====
    constexpr uint32_t idesc = make_idesc_i8(PAIR ? 2 * TC_BM : TC_BM, TC_BN, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    bool ready = false;
    for (int t = unit0; t < num_tiles && rank == 0; t += nunits) {
      mbar_wait_warp(&tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * TC_BN;
      const int kb0 = tiles[t].kb0, kb1 = tiles[t].kb1 < 0 ? num_kblocks : tiles[t].kb1;
      for (int kb = kb0; kb < kb1; kb++) {
        if (!ready) mbar_wait_warp(&full[stage], phase);
        tc_fence_after();
        // a try_wait costs ~90 cycles even on a completed phase: probe the NEXT stage now
        // (non-blocking) and consume the answer after this stage's MMAs are issued
        const int stage_n = stage + 1 == TC_STAGES ? 0 : stage + 1;
        const uint32_t phase_n = stage + 1 == TC_STAGES ? phase ^ 1 : phase;
        const bool poll = mbar_test_wait(&full[stage_n], phase_n);
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * TC_STAGE_BYTES);
          const uint64_t adesc = make_desc_k_sw128(sa);
          const uint64_t bdesc = make_desc_k_sw128(sa + TC_A_BYTES);
#pragma unroll
          for (int k4 = 0; k4 < TC_BK / 32; k4++) {  // +32 bytes along K = +2 in the >>4 address field
            if (PAIR) mma_i8_ss_pair(d_tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc, (kb > kb0 || k4) ? 1u : 0u);
            else mma_i8_ss(d_tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc, (kb > kb0 || k4) ? 1u : 0u);
          }
          if (PAIR) {  // the stage is free, and the accumulator ready, in BOTH CTAs
            tc_commit_pair(&empty[stage], 3);
            if (kb == kb1 - 1) tc_commit_pair(&tfull[acc], 3);
          } else {
            tc_commit(&empty[stage]);
            if (kb == kb1 - 1) tc_commit(&tfull[acc]);
          }
        }
        __syncwarp();
        ready = __all_sync(0xffffffffu, poll);
        stage = stage_n;
        phase = phase_n;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quadrant = warp % 4 =====
    const int quad = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    const CpAffine aff = cp_affine_setup(fin.n, fin.affine, fin.do_scale, fin.scal);
    double *stg = reinterpret_cast<double *>(smem + TC_STAGES * TC_STAGE_BYTES + 256) + (warp - 2) * TC_STG;
    for (int t = unit0; t < num_tiles; t += nunits) {
      TileCoord tc = tiles[t];
      if (PAIR) {  // this CTA's half of the pair
        tc.mi += rank * mstride;
        tc.mo += rank;
        if (tc.slot >= 0) tc.slot = 2 * tc.slot + rank;
      }
      // second half of a strided pair whose tile lies wholly left of the diagonal: computed by the
      // 256-row MMA, not part of the result
      const bool dead = PAIR && (long)TC_BN * (tc.nj + 1) <= (long)TC_BM * tc.mi;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const long row = (long)tc.mo * TC_BM + quad * 32 + lane;
      int32_t *crow = C32 + row * ld32 + (long)tc.nj * TC_BN;
      const uint32_t taddr = tmem_base + acc * TC_BN + ((uint32_t)(quad * 32) << 16);
      // global row of the result this thread owns (panels are compact in the row tiles)
      const long gi = fin.kind == 2 ? ((long)tc.mo * fin.nparts + fin.part) * TC_BM + quad * 32 + lane
                                    : row;
      const bool iok = fin.kind && gi < fin.n;
      const long long ai = (fin.affine && iok) ? __double2ll_rn(fin.m[gi]) * aff.n : 0;
#pragma unroll 1
      for (int c = 0; c < TC_BN / 32 && !dead; c++) {
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
        const long long blane =
            (fin.affine && j0 + lane < fin.n) ? __double2ll_rn(__ldg(fin.m + j0 + lane)) * aff.n - aff.total : 0;
        long long bjv[32];
#pragma unroll
        for (int q = 0; q < 32; q++) bjv[q] = __shfl_sync(0xffffffffu, blane, q);
        // Stores whose consecutive addresses belong to ONE thread (the mirrored element of the full
        // matrix, the row-major panel of a partition) would cost 32 L1 wavefronts per instruction
        // and starve the tensor core's operand reads, which share the L1 data pipe: the 32 x 32
        // block goes through a per-warp staging tile and leaves with lanes = consecutive addresses.
        const long rowbase = gi - lane;  // global row of lane 0's thread
        double *col = fin.out + gi + j0 * fin.n;        // kind 1: (i, j0 + q), lanes = consecutive i
#pragma unroll
        for (int q = 0; q < 32; q++) {
          const double v = fin.affine ? cp_affine((int)r[q], ai, bjv[q], aff) : (double)(int)r[q];
          stg[lane * 33 + q] = v;
          if (fin.kind == 1 && iok && j0 + q < fin.n && gi <= j0 + q) col[(long)q * fin.n] = v;
        }
        __syncwarp();
        const long j = j0 + lane;                       // this lane's column of the 32 x 32 block
        if (fin.kind == 1) {                            // mirrored element (j, i): exactly symmetric
#pragma unroll 8
          for (int i = 0; i < 32; i++) {
            const long gr = rowbase + i;
            if (j < fin.n && gr < fin.n && gr <= j) fin.out[j + gr * fin.n] = stg[i * 33 + lane];
          }
        } else {                                        // row-major panel of a partition
          double *pblock = fin.out + (row - lane) * fin.n + j;
#pragma unroll 8
          for (int i = 0; i < 32; i++)
            if (j < fin.n && rowbase + i < fin.n) pblock[(long)i * fin.n] = stg[i * 33 + lane];
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR && rank) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty[acc]), 0));  // the issuer waits on rank 0's
        else mbar_arrive(&tempty[acc]);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the leader's MMAs read the peer's smem and write its TMEM
  if (warp == 0) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
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
  if ((long)TC_BN * (tc.nj + 1) <= (long)TC_BM * tc.mi) return;  // dead half of a strided pair
  const CpAffine aff = cp_affine_setup(fin.n, fin.affine, fin.do_scale, fin.scal);
  const int *src = scratch + (long)blockIdx.x * TC_BM * TC_BN;
  for (int e = threadIdx.x; e < TC_BM * TC_BN; e += 256) {
    // kind 1 stores column-major (consecutive threads = consecutive rows), kind 2 row-major
    const int r = fin.kind == 1 ? e % TC_BM : e / TC_BN, c = fin.kind == 1 ? e / TC_BM : e % TC_BN;
    const long lrow = (long)tc.mo * TC_BM + r;
    const long gi = fin.kind == 2 ? ((long)tc.mo * fin.nparts + fin.part) * TC_BM + r : lrow;
    const long j = (long)tc.nj * TC_BN + c;
    if (gi >= fin.n || j >= fin.n) continue;
    const int cnt = src[r * TC_BN + c];
    const double v = fin.affine ? cp_affine_m(cnt, fin.m[gi], fin.m[j], aff) : (double)cnt;
    if (fin.kind == 1) {
      if (gi > j) continue;
      fin.out[gi + j * fin.n] = v;
      fin.out[j + gi * fin.n] = v;
    } else {
      fin.out[lrow * fin.n + j] = v;
    }
  }
}

// split_kblocks > 0: K blocks per tile of a single-slab product whose last partial round may be
// cut into K pieces (see TileCoord); *num_tail = tiles cut, their coordinates follow the work list
// at (*d_tiles)[*num_tiles ...].
int build_tile_list(long rows, int part, int nparts, TileCoord **d_tiles, int *num_tiles,
                    cudaStream_t stream, int split_kblocks, int grid_max, int *num_tail,
                    int *real_tiles, int pair) {
  const int Mt = (int)ceil_div(rows, TC_BM), Nt = (int)ceil_div(rows, TC_BN);
  std::vector<TileCoord> v;
  if (pair) {
    // work items = pairs of this partition's consecutive row tiles (mi, mi + nparts) against column
    // tile nj, listed when the first half touches the upper triangle.  nparts == 1: both halves of
    // (2 Mi, 2 Mi + 1) are upper-triangle tiles for nj >= Mi.  nparts > 1: the second half may lie
    // wholly left of the diagonal (at most nparts / 2 column tiles per pair; the epilogue drops it)
    // or past the last row tile (odd tile count; TMA zero-fills, the epilogue's row guard drops it).
    for (int sn = 0; sn < Nt; sn += 8)
      for (int sm = 0; sm < Mt; sm += 16 * nparts)
        for (int nj = sn; nj < sn + 8 && nj < Nt; nj++)
          for (int mi = sm + part; mi < sm + 16 * nparts && mi < Mt; mi += 2 * nparts)
            if ((long)TC_BN * (nj + 1) > (long)TC_BM * mi) v.push_back({mi, nj, mi / nparts, 0, -1, -1});
  } else {
    for (int sn = 0; sn < Nt; sn += 8)
      for (int sm = 0; sm < Mt; sm += 16 * nparts)
        for (int nj = sn; nj < sn + 8 && nj < Nt; nj++)
          for (int mi = sm + part; mi < sm + 16 * nparts && mi < Mt; mi += nparts)
            if ((long)TC_BN * (nj + 1) > (long)TC_BM * mi) v.push_back({mi, nj, mi / nparts, 0, -1, -1});
  }
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
      ntail = pair ? 2 * rem : rem;                      // scratch tiles of 128 x 256
      *num_tiles = (int)v.size();
      for (const TileCoord &c : tail) {                  // finish list: one entry per scratch tile
        v.push_back(c);
        if (pair) v.push_back({c.mi + nparts, c.nj, c.mo + 1, 0, -1, -1});
      }
    }
  }
  if (!ntail) *num_tiles = T;
  if (num_tail) *num_tail = ntail;
  if (real_tiles) *real_tiles = pair ? 2 * T : T;
  if (v.empty()) v.push_back({0, 0, 0, 0, -1, -1});  // keep the allocation non-empty; the launch is skipped
  MB_CUDA(pool_alloc(d_tiles, sizeof(TileCoord) * v.size(), stream));
  MB_CUDA(cudaMemcpyAsync(*d_tiles, v.data(), sizeof(TileCoord) * v.size(), cudaMemcpyHostToDevice,
                          stream));
  MB_CUDA(cudaStreamSynchronize(stream));  // v goes out of scope
  return 0;
}

// Tile list of a BAND of consecutive row tiles [tile_begin, tile_end) against all column tiles at
// or right of the diagonal (the out-of-core host path, crossprod_host.cu: a band's rows are
// consecutive rows of the result, so its panels leave for the host as large contiguous copies).
// Output row tile mo = mi - tile_begin.  pair: CTA pairs over (mi, mi + 1); tile_begin must be even.
int build_band_tile_list(long rows, int tile_begin, int tile_end, TileCoord **d_tiles, int *num_tiles,
                         cudaStream_t stream, int *real_tiles, int pair) {
  const int Mt = (int)ceil_div(rows, TC_BM), Nt = (int)ceil_div(rows, TC_BN);
  if (tile_end > Mt) tile_end = Mt;
  std::vector<TileCoord> v;
  const int step = pair ? 2 : 1;
  for (int sn = 0; sn < Nt; sn += 8)
    for (int sm = tile_begin; sm < tile_end; sm += 16)
      for (int nj = sn; nj < sn + 8 && nj < Nt; nj++)
        for (int mi = sm; mi < sm + 16 && mi < tile_end; mi += step)
          if ((long)TC_BN * (nj + 1) > (long)TC_BM * mi) v.push_back({mi, nj, mi - tile_begin, 0, -1, -1});
  *num_tiles = (int)v.size();
  if (real_tiles) *real_tiles = pair ? 2 * (int)v.size() : (int)v.size();
  if (v.empty()) v.push_back({0, 0, 0, 0, -1, -1});
  MB_CUDA(pool_alloc(d_tiles, sizeof(TileCoord) * v.size(), stream));
  MB_CUDA(cudaMemcpyAsync(*d_tiles, v.data(), sizeof(TileCoord) * v.size(), cudaMemcpyHostToDevice, stream));
  MB_CUDA(cudaStreamSynchronize(stream));  // v goes out of scope
  return 0;
}

// Work items one launch keeps in flight: CTAs (one per SM) or CTA pairs (clusters of 2).
int crossprod_tc_units(int sm_count, int pair) {
  if (!pair) return sm_count;
  // cached per device: the shared-memory opt-in and the cluster occupancy are device properties
  // (one process may drive several devices, multi.cu / crossprod_host.cu)
  static std::atomic<int> cache[64];
  static std::atomic<bool> cache_init{false};
  static std::mutex mu;
  int device = 0;
  cudaGetDevice(&device);
  const bool cached_dev = device >= 0 && device < 64;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!cache_init.load()) {
      for (int i = 0; i < 64; i++) cache[i].store(-1);
      cache_init.store(true);
    }
  }
  int clusters = cached_dev ? cache[device].load() : -1;
  if (clusters < 0) {
    cudaFuncSetAttribute(crossprod_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * sm_count);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = TC_SMEM;
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeClusterDimension;
    at.val.clusterDim.x = 2;
    at.val.clusterDim.y = 1;
    at.val.clusterDim.z = 1;
    cfg.attrs = &at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, crossprod_tc_kernel<true>, &cfg) != cudaSuccess || n < 1) {
      cudaGetLastError();
      n = 0;
    }
    clusters = n < sm_count / 2 ? n : sm_count / 2;
    if (cached_dev) cache[device].store(clusters);
  }
  return clusters;
}

int crossprod_tc_launch(const int8_t *X8, long rows_pad, long kcols, long pitch,
                        const TileCoord *d_tiles,
                        int num_tiles, int32_t *C32, long ld32, int accumulate, int sm_count,
                        cudaStream_t stream, const CpFinish *finish, int num_tail, int pair) {
  // pair: 0 = one CTA per 128 x 256 tile; s >= 1 = CTA pairs over row tiles (mi, mi + s)
  CUtensorMap ta, tb;
  MB_CHECK(make_tmap_u8_2d(&ta, X8, kcols, rows_pad, pitch, TC_BK, TC_BM, 1));
  MB_CHECK(make_tmap_u8_2d(&tb, X8, kcols, rows_pad, pitch, TC_BK, TC_BN, 1));
  if (num_tiles <= 0) return 0;
  const int units = crossprod_tc_units(sm_count, pair);
  if (units < 1) return fail(-12, __func__, __LINE__, "no co-resident CTA pairs on this device");
  const int grid = (num_tiles < units ? num_tiles : units) * (pair ? 2 : 1);
  CpFinish fin = finish ? *finish : CpFinish();
  int *scratch = nullptr;
  if (num_tail > 0) {
    if (!finish || !finish->kind) return fail(-2, __func__, __LINE__, "split tail needs a finishing epilogue");
    const size_t bytes = sizeof(int) * (size_t)num_tail * TC_BM * TC_BN;
    MB_CUDA(pool_alloc(&scratch, bytes, stream));
    MB_CUDA(cudaMemsetAsync(scratch, 0, bytes, stream));
    fin.tail = scratch;
  }
  if (pair) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = TC_SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeClusterDimension;
    at.val.clusterDim.x = 2;
    at.val.clusterDim.y = 1;
    at.val.clusterDim.z = 1;
    cfg.attrs = &at;
    cfg.numAttrs = 1;
    MB_CUDA(cudaLaunchKernelEx(&cfg, crossprod_tc_kernel<true>, ta, tb, d_tiles, num_tiles,
                               (int)(kcols / TC_BK), C32, ld32, accumulate, fin, pair));
  } else {
    MB_CUDA(cudaFuncSetAttribute(crossprod_tc_kernel<false>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
    crossprod_tc_kernel<false><<<grid, TC_THREADS, TC_SMEM, stream>>>(ta, tb, d_tiles, num_tiles,
                                                                      (int)(kcols / TC_BK), C32, ld32,
                                                                      accumulate, fin, 1);
  }
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
