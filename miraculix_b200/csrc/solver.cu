// solver.cu -- device-resident conjugate-gradient solve of (Zc Zc^T + lambda I) x = b on a
// compressed handle (BASELINE configs[3]).
//
// Replaces the host loop of the reference's examples/iterative_solver/grm_solve_cg.jl:109-133
// (GRM_vec :75-85 = dgemm_compressed 'T' then 'N' with one column, every iteration through host
// memory, BLAS-1 in Julia).  Here the packed matrix AND all vectors stay in HBM: one iteration is
//   t  = Zc^T p            dgemm_compressed 'T', n = 1     (local to a SNP shard)
//   Gp = Zc t              dgemm_compressed 'N', n = 1     (+ caller's all-reduce when sharded)
//   cg_pgp_kernel          Gp += lambda p ; pGp = p.Gp ; alpha = rr / pGp
//   cg_update_kernel       x += alpha p ; r -= alpha Gp ; rr' = r.r ; beta = rr'/rr
//   cg_direction_kernel    p = r + beta p
// with the same recurrences, in the same order, as the reference script.  Reductions are
// deterministic (fixed partial grid, last block sums the partials in order), so every rank of a
// SNP-sharded solve computes bit-identical scalars from the all-reduced Gp.
#include <math.h>
#include "../../include/miraculix_b200.h"
#include "common.cuh"
#include "handle.h"
#include "multi.h"

namespace mb200 {

constexpr int CG_BLOCKS = 296;  // 2 per SM
constexpr int CG_THREADS = 256;

// scalars on the device: [0] rr = r.r, [1] pGp, [2] alpha, [3] beta, [4] rr of the previous iterate
struct CgScalars {
  double rr, pgp, alpha, beta, rr_old;
};

__device__ __forceinline__ double block_sum(double v, double *red) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < CG_THREADS / 32; i++) t += red[i];
  return t;  // valid in thread 0
}

// Returns true in thread 0 of the LAST block to finish, with *total = sum of all partials in
// block order (deterministic); resets the ticket for the next launch.
__device__ __forceinline__ bool grid_sum(double part, double *partials, unsigned *ticket,
                                         double *total) {
  __shared__ bool last;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = part;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last || threadIdx.x != 0) return false;
  __threadfence();
  double t = 0.0;
  for (unsigned i = 0; i < gridDim.x; i++) t += reinterpret_cast<volatile double *>(partials)[i];
  *total = t;
  *ticket = 0;
  return true;
}

// r = b - Gx (Gx already holds Zc Zc^T x; the lambda x term is added here), p = r, rr = r.r
__global__ void __launch_bounds__(CG_THREADS)
cg_init_kernel(long n, const double *__restrict__ b, const double *__restrict__ x,
               const double *__restrict__ Gx, double lambda, double *__restrict__ r,
               double *__restrict__ p, CgScalars *sc, double *partials, unsigned *ticket,
               double *host_rr) {
  __shared__ double red[CG_THREADS / 32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += (long)gridDim.x * CG_THREADS) {
    const double ri = b[i] - (Gx[i] + lambda * x[i]);
    r[i] = ri;
    p[i] = ri;
    acc += ri * ri;
  }
  const double part = block_sum(acc, red);
  double tot;
  if (grid_sum(part, partials, ticket, &tot)) {
    sc->rr = tot;
    sc->rr_old = tot;
    *host_rr = tot;
  }
}

__global__ void __launch_bounds__(CG_THREADS)
cg_pgp_kernel(long n, const double *__restrict__ p, double *__restrict__ Gp, double lambda,
              CgScalars *sc, double *partials, unsigned *ticket) {
  __shared__ double red[CG_THREADS / 32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += (long)gridDim.x * CG_THREADS) {
    const double g = Gp[i] + lambda * p[i];
    Gp[i] = g;
    acc += p[i] * g;
  }
  const double part = block_sum(acc, red);
  double tot;
  if (grid_sum(part, partials, ticket, &tot)) {
    sc->pgp = tot;
    sc->alpha = sc->rr / tot;  // grm_solve_cg.jl:123
  }
}

__global__ void __launch_bounds__(CG_THREADS)
cg_update_kernel(long n, const double *__restrict__ p, const double *__restrict__ Gp,
                 double *__restrict__ x, double *__restrict__ r, CgScalars *sc, double *partials,
                 unsigned *ticket, double *host_rr) {
  __shared__ double red[CG_THREADS / 32];
  const double alpha = sc->alpha;
  double acc = 0.0;
  for (long i = (long)blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += (long)gridDim.x * CG_THREADS) {
    x[i] += alpha * p[i];          // :126
    const double ri = r[i] - alpha * Gp[i];  // :127
    r[i] = ri;
    acc += ri * ri;
  }
  const double part = block_sum(acc, red);
  double tot;
  if (grid_sum(part, partials, ticket, &tot)) {
    sc->rr_old = sc->rr;
    sc->beta = tot / sc->rr;  // :128
    sc->rr = tot;
    *host_rr = tot;
  }
}

__global__ void __launch_bounds__(CG_THREADS)
cg_direction_kernel(long n, const double *__restrict__ r, double *__restrict__ p,
                    const CgScalars *sc) {
  const double beta = sc->beta;
  for (long i = (long)blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += (long)gridDim.x * CG_THREADS)
    p[i] = r[i] + beta * p[i];  // :129
}

struct CgWork {
  double *t, *Gp, *r, *p, *partials;
  CgScalars *sc;
  unsigned *ticket;
};

static int cg_workspace(Handle *h, CgWork *w) {
  const size_t need = sizeof(double) * ((size_t)h->snps + 3 * (size_t)h->indiv + CG_BLOCKS + 16);
  if (!h->d_cg || h->cg_bytes < need) {
    if (h->d_cg) MB_CUDA(cudaFree(h->d_cg));
    h->d_cg = nullptr;
    MB_CUDA(cudaMalloc(&h->d_cg, need));
    h->cg_bytes = need;
  }
  if (!h->h_cg) MB_CUDA(cudaMallocHost(&h->h_cg, sizeof(double) * 4));
  double *base = reinterpret_cast<double *>(h->d_cg);
  w->t = base;
  w->Gp = w->t + h->snps;
  w->r = w->Gp + h->indiv;
  w->p = w->r + h->indiv;
  w->partials = w->p + h->indiv;
  w->sc = reinterpret_cast<CgScalars *>(w->partials + CG_BLOCKS);
  w->ticket = reinterpret_cast<unsigned *>(w->sc + 1);
  MB_CUDA(cudaMemsetAsync(w->sc, 0, sizeof(CgScalars) + 2 * sizeof(double), h->stream));
  return 0;
}

// Gv = Zc (Zc^T v) on this handle's SNP range; all-reduced over the shards by the caller's hook
static int grm_vec(Handle *h, const double *v, const CgWork &w, mb200_allreduce_fn ar, void *user) {
  MB_CHECK(dgemm_device(h, 1, 1, v, h->indiv, w.t, h->snps));
  MB_CHECK(dgemm_device(h, 0, 1, w.t, h->snps, w.Gp, h->indiv));
  if (ar) {
    const int rc = ar(w.Gp, h->indiv, (void *)h->stream, user);
    if (rc) return fail(rc, __func__, __LINE__, "all-reduce callback returned %d", rc);
  }
  return 0;
}

int cg_solve_device(Handle *h, const double *d_b, double *d_x, double lambda, double tol,
                    int max_iter, mb200_allreduce_fn ar, void *user, int *iters, double *resid) {
  CgWork w;
  MB_CHECK(cg_workspace(h, &w));
  const long n = h->indiv;
  double *host_rr = reinterpret_cast<double *>(h->h_cg);
  // host_rr is pinned and mapped: the kernels' last block writes it, the host reads after a sync
  double *dev_rr = nullptr;
  MB_CUDA(cudaHostGetDevicePointer(&dev_rr, host_rr, 0));

  MB_CHECK(grm_vec(h, d_x, w, ar, user));  // Gx0, :113
  cg_init_kernel<<<CG_BLOCKS, CG_THREADS, 0, h->stream>>>(n, d_b, d_x, w.Gp, lambda, w.r, w.p, w.sc,
                                                          w.partials, w.ticket, dev_rr);
  count_launch();
  MB_LAUNCH_CHECK();
  MB_CUDA(cudaStreamSynchronize(h->stream));
  int it = 0;
  double rr = *reinterpret_cast<volatile double *>(host_rr);
  for (; it < max_iter; it++) {
    if (!(sqrt(rr) >= tol)) break;  // :118-121 (also leaves on NaN)
    MB_CHECK(grm_vec(h, w.p, w, ar, user));
    cg_pgp_kernel<<<CG_BLOCKS, CG_THREADS, 0, h->stream>>>(n, w.p, w.Gp, lambda, w.sc, w.partials,
                                                           w.ticket);
    cg_update_kernel<<<CG_BLOCKS, CG_THREADS, 0, h->stream>>>(n, w.p, w.Gp, d_x, w.r, w.sc,
                                                              w.partials, w.ticket, dev_rr);
    cg_direction_kernel<<<CG_BLOCKS, CG_THREADS, 0, h->stream>>>(n, w.r, w.p, w.sc);
    count_launch(3);
    MB_LAUNCH_CHECK();
    MB_CUDA(cudaStreamSynchronize(h->stream));
    rr = *reinterpret_cast<volatile double *>(host_rr);
  }
  if (iters) *iters = it;
  if (resid) *resid = sqrt(rr);
  return 0;
}

}  // namespace mb200

using namespace mb200;

extern "C" {

int mb200_cg_solve_device(void *compressed, const double *d_b, double *d_x, double lambda,
                          double tol, int max_iter, mb200_allreduce_fn allreduce, void *user,
                          int *iters, double *resid_norm) {
  clear_status();
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  if (!d_b || !d_x || max_iter < 0)
    return fail(-2, __func__, __LINE__, "bad arguments (b=%p x=%p max_iter=%d)", (const void *)d_b,
                (void *)d_x, max_iter);
  return cg_solve_device(h, d_b, d_x, lambda, tol, max_iter, allreduce, user, iters, resid_norm);
}

int mb200_cg_solve(void *compressed, const double *b, double *x, double lambda, double tol,
                   int max_iter, int *iters, double *resid_norm) {
  MB_RANGE();
  clear_status();
  if (MultiHandle *m = as_multi(compressed)) {
    if (!b || !x || max_iter < 0) return fail(-2, __func__, __LINE__, "bad arguments");
    return multi_cg_solve(m, b, x, lambda, tol, max_iter, iters, resid_norm);
  }
  Handle *h = as_handle(compressed);
  if (!h) return status().code;
  if (!b || !x || max_iter < 0) return fail(-2, __func__, __LINE__, "bad arguments");
  const size_t bytes = sizeof(double) * (size_t)h->indiv;
  double *d = nullptr;
  MB_CUDA(cudaMalloc(&d, 2 * bytes));
  int rc = 0;
  cudaError_t e = cudaMemcpyAsync(d, b, bytes, cudaMemcpyHostToDevice, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d + h->indiv, x, bytes, cudaMemcpyHostToDevice, h->stream);
  if (e != cudaSuccess) rc = fail((int)e, __func__, __LINE__, "H2D: %s", cudaGetErrorString(e));
  if (!rc) rc = cg_solve_device(h, d, d + h->indiv, lambda, tol, max_iter, nullptr, nullptr, iters,
                                resid_norm);
  if (!rc) {
    e = cudaMemcpyAsync(x, d + h->indiv, bytes, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) rc = fail((int)e, __func__, __LINE__, "D2H: %s", cudaGetErrorString(e));
  }
  cudaFree(d);
  return rc;
}

}  // extern "C"
