// pcg.cu — block-Jacobi preconditioned conjugate gradients on the tiled reduced camera system.
//
// Second implementation of CalculateGn (reference src/BundleAdjuster.cpp:748-833: solve S delta_p = rhs_sc): north_star asks
// for "a dense blocked Cholesky on the FP64 tensor pipe when the pose count is large, and a block-Jacobi PCG otherwise".
// Selected with options.reduced_solver (1 = always, 2 = automatically for small systems).
//
// S is the upper block-tile storage of ba_gpu.h (tiles of T x T doubles, column-major, permuted tile ids, diagonal tiles
// hold their upper triangle).  Only tiles of the symbolic pattern are visited (fill tiles of the factorisation are zero).
//   y = S x      one CTA per tile row, one thread per row of the tile; the row's tiles (its own row of the upper
//                storage and, transposed, the tiles of its column) are listed once per Solve() on the host, in a fixed
//                order: no atomics, bit-reproducible
//   M^-1         inverse of the PoseSize x PoseSize diagonal block of every pose (Cholesky, one thread per pose)
//   the loop     runs on the device: alpha and beta are formed by every thread from scalars that fixed-order grid
//                reductions left in device memory; the host looks at the `done` flag once per chunk of iterations.
// A non-positive p^T S p or a non-positive pivot of a diagonal block reports the system as not positive definite
// (FactorizationError, like a failed LDL^T).
#include <algorithm>
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"

namespace bagpu {

// scalars of the running solve
struct PcgScalars {
  double rz[2];     // r.z of iteration k at [k & 1]
  double pAp;
  double rr;        // r.r
  double r0;        // |rhs|^2
  double tmp[2];    // (r.z, r.r) of the update in flight
  int done;         // 1 = converged or out of iterations, 2 = breakdown (not positive definite)
  int iters;
};

// ---- block-Jacobi preconditioner: M^-1 = inverse of the D x D diagonal blocks ----------------------------------------
template <int D>
__global__ void k_pcg_precond(const double* __restrict__ S, const int* __restrict__ tile_map, int nT, int T, double* __restrict__ Minv,
                              PcgScalars* sc) {
  const int pose = blockIdx.x * blockDim.x + threadIdx.x;   // position in the permuted, padded order
  if (pose >= nT * kTilePoses) return;
  const int t = pose / kTilePoses, o = (pose % kTilePoses) * D;
  const double* tile = S + (size_t)tile_map[(size_t)t * nT + t] * T * T;
  double A[D * D], Li[D * D];
  for (int c = 0; c < D; ++c) for (int r = 0; r <= c; ++r) { const double v = tile[(size_t)(o + c) * T + (o + r)]; A[r * D + c] = v; A[c * D + r] = v; }
  // Cholesky A = L L^T (lower, in place), then L^-1, then A^-1 = L^-T L^-1
  bool ok = true;
  for (int j = 0; j < D; ++j) {
    double d = A[j * D + j];
    for (int k = 0; k < j; ++k) d -= A[j * D + k] * A[j * D + k];
    if (!(d > 0.0)) { ok = false; d = 1.0; }
    const double l = sqrt(d);
    A[j * D + j] = l;
    for (int i = j + 1; i < D; ++i) {
      double s = A[i * D + j];
      for (int k = 0; k < j; ++k) s -= A[i * D + k] * A[j * D + k];
      A[i * D + j] = s / l;
    }
  }
  for (int i = 0; i < D * D; ++i) Li[i] = 0.0;
  for (int j = 0; j < D; ++j) {
    Li[j * D + j] = 1.0 / A[j * D + j];
    for (int i = j + 1; i < D; ++i) {
      double s = 0.0;
      for (int k = j; k < i; ++k) s -= A[i * D + k] * Li[k * D + j];
      Li[i * D + j] = s / A[i * D + i];
    }
  }
  double* out = Minv + (size_t)pose * D * D;
  for (int r = 0; r < D; ++r) for (int c = 0; c < D; ++c) {
    double s = 0.0;
    for (int k = (r > c ? r : c); k < D; ++k) s += Li[k * D + r] * Li[k * D + c];
    out[r * D + c] = s;
  }
  if (!ok) sc->done = 2;
}

// ---- y = S x ------------------------------------------------------------------------------------------------------------
// entries of tile row a: (slot, other tile, kind) with kind 0 = tile (a, other) as stored, 1 = tile (other, a) transposed,
// 2 = the diagonal tile (upper triangle stored)
__global__ void __launch_bounds__(128) k_pcg_spmv(const double* __restrict__ S, const int* __restrict__ row_ptr, const int3* __restrict__ ent, int T,
                                                  const double* __restrict__ x, double* __restrict__ y, const PcgScalars* sc) {
  if (sc->done) return;
  extern __shared__ double xs[];
  const int a = blockIdx.x, i = threadIdx.x;
  double acc = 0.0;
  for (int e = row_ptr[a]; e < row_ptr[a + 1]; ++e) {
    const int3 en = ent[e];
    __syncthreads();
    if (i < T) xs[i] = x[(size_t)en.y * T + i];
    __syncthreads();
    if (i >= T) continue;
    const double* tile = S + (size_t)en.x * T * T;
    if (en.z == 0) {
      for (int c = 0; c < T; ++c) acc += tile[(size_t)c * T + i] * xs[c];
    } else if (en.z == 1) {
      const double* col = tile + (size_t)i * T;
      for (int r = 0; r < T; ++r) acc += col[r] * xs[r];
    } else {
      const double* col = tile + (size_t)i * T;
      for (int r = 0; r < i; ++r) acc += col[r] * xs[r];
      for (int c = i; c < T; ++c) acc += tile[(size_t)c * T + i] * xs[c];
    }
  }
  if (i < T) y[(size_t)a * T + i] = acc;
}

// ---- vector kernels (one CTA per tile, T threads: the preconditioner needs whole pose blocks) ---------------------------
// z = M^-1 r for the thread's element
template <int D>
BA_D double precond_apply(const double* __restrict__ Minv, const double* rs, int tile, int i) {
  const int pl = i / D, k = i - pl * D;
  const double* M = Minv + ((size_t)tile * kTilePoses + pl) * D * D + (size_t)k * D;
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < D; ++c) s += M[c] * rs[pl * D + c];
  return s;
}

// start: x = 0, r = b, z = M^-1 r, p = z, tmp = (r.z, r.r)
template <int D>
__global__ void __launch_bounds__(128) k_pcg_init(const double* __restrict__ b, const double* __restrict__ Minv, double* x, double* r, double* z,
                                                  double* p, PcgScalars* sc, double* partial, unsigned int* counter) {
  __shared__ double rs[kTilePoses * D];
  const int T = kTilePoses * D, i = threadIdx.x, t = blockIdx.x;
  double v[2] = {0.0, 0.0};
  double ri = 0.0;
  if (i < T) { ri = b[(size_t)t * T + i]; rs[i] = ri; }
  __syncthreads();
  if (i < T) {
    const double zi = precond_apply<D>(Minv, rs, t, i);
    const size_t g = (size_t)t * T + i;
    x[g] = 0.0; r[g] = ri; z[g] = zi; p[g] = zi;
    v[0] = ri * zi; v[1] = ri * ri;
  }
  grid_sum_fixed<2>(v, partial, counter, &sc->tmp[0], false);
}
__global__ void k_pcg_init_scalars(PcgScalars* sc) {
  sc->rz[0] = sc->tmp[0]; sc->rz[1] = 0.0;
  sc->r0 = sc->tmp[1]; sc->rr = sc->tmp[1]; sc->pAp = 0.0; sc->iters = 0;
  if (sc->done != 2) sc->done = (sc->tmp[1] == 0.0) ? 1 : 0;
}

// pAp = p . Ap
__global__ void __launch_bounds__(256) k_pcg_dot(const double* __restrict__ a, const double* __restrict__ b, size_t n, PcgScalars* sc, double* partial,
                                                 unsigned int* counter) {
  double v[1] = {0.0};
  if (!sc->done) for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) v[0] += a[i] * b[i];
  grid_sum_fixed<1>(v, partial, counter, &sc->pAp, false);
}

// x += alpha p, r -= alpha Ap, z = M^-1 r, tmp = (r.z, r.r)
template <int D>
__global__ void __launch_bounds__(128) k_pcg_update(const double* __restrict__ Minv, const double* __restrict__ p, const double* __restrict__ Ap,
                                                    double* x, double* r, double* z, PcgScalars* sc, int k, double* partial, unsigned int* counter) {
  __shared__ double rs[kTilePoses * D];
  const int T = kTilePoses * D, i = threadIdx.x, t = blockIdx.x;
  const double pAp = sc->pAp;
  const bool live = !sc->done && pAp > 0.0;
  const double alpha = live ? sc->rz[k & 1] / pAp : 0.0;
  double v[2] = {0.0, 0.0};
  const size_t g = (size_t)t * T + i;
  double ri = 0.0;
  if (i < T) {
    ri = r[g];
    if (live) { x[g] += alpha * p[g]; ri -= alpha * Ap[g]; r[g] = ri; }
    rs[i] = ri;
  }
  __syncthreads();
  if (i < T) {
    const double zi = precond_apply<D>(Minv, rs, t, i);
    if (live) z[g] = zi;
    v[0] = ri * zi; v[1] = ri * ri;
  }
  grid_sum_fixed<2>(v, partial, counter, &sc->tmp[0], false);
}

// bookkeeping of iteration k (one thread): breakdown, convergence, iteration count
__global__ void k_pcg_book(PcgScalars* sc, int k, double tol2, int max_iter) {
  if (sc->done) return;
  if (!(sc->pAp > 0.0)) { sc->done = 2; return; }
  sc->rz[(k + 1) & 1] = sc->tmp[0];
  sc->rr = sc->tmp[1];
  sc->iters = k + 1;
  if (sc->tmp[1] <= tol2 * sc->r0 || k + 1 >= max_iter) sc->done = 1;
}

// p = z + beta p
__global__ void __launch_bounds__(256) k_pcg_next(const double* __restrict__ z, double* p, size_t n, const PcgScalars* sc, int k) {
  if (sc->done) return;
  const double beta = sc->rz[(k + 1) & 1] / sc->rz[k & 1];
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = z[i] + beta * p[i];
}

// ---- host driver --------------------------------------------------------------------------------------------------------
size_t pcg_work_doubles(int nT, int T, int D) { return 5 * (size_t)nT * T + (size_t)nT * kTilePoses * D * D + 16; }

// Solves S x = b (b, x: permuted tile order, padded to nT * T).  work: pcg_work_doubles() doubles.  Returns 0 = converged,
// 1 = out of iterations, 2 = not positive definite; < 0 = CUDA error.  *iters_out = iterations done.
int launch_pcg_solve(int D, const double* S, const int* tile_map, const int* row_ptr, const int3* ent, int nT, const double* b, double* x_out,
                     double* work, double* red_partial, unsigned int* red_counter, double tol, int max_iter, int* iters_out, uint64_t* launches,
                     cudaStream_t st) {
  const int T = kTilePoses * D;
  const size_t n = (size_t)nT * T;
  double* x = work; double* r = x + n; double* z = r + n; double* p = z + n; double* Ap = p + n;
  double* Minv = Ap + n;
  PcgScalars* sc = reinterpret_cast<PcgScalars*>(Minv + (size_t)nT * kTilePoses * D * D);
  cudaMemsetAsync(sc, 0, sizeof(PcgScalars), st);
  const int np = nT * kTilePoses;
  const int gv = (int)std::min<size_t>((n + 255) / 256, (size_t)kRedMaxGrid);
#define PCG_D(CALL) do { if (D == 6) { CALL(6); } else if (D == 9) { CALL(9); } else { CALL(15); } } while (0)
#define PRE(DD) k_pcg_precond<DD><<<(np + 63) / 64, 64, 0, st>>>(S, tile_map, nT, T, Minv, sc)
  PCG_D(PRE);
#define INI(DD) k_pcg_init<DD><<<nT, 128, 0, st>>>(b, Minv, x, r, z, p, sc, red_partial, red_counter)
  PCG_D(INI);
  k_pcg_init_scalars<<<1, 1, 0, st>>>(sc);
  *launches += 3;
  const double tol2 = tol * tol;
  const int chunk = 16;
  int k = 0, done = 0, iters = 0;
  while (k < max_iter) {
    for (int c = 0; c < chunk && k < max_iter; ++c, ++k) {
      k_pcg_spmv<<<nT, 128, sizeof(double) * T, st>>>(S, row_ptr, ent, T, p, Ap, sc);
      k_pcg_dot<<<gv, 256, 0, st>>>(p, Ap, n, sc, red_partial, red_counter);
#define UPD(DD) k_pcg_update<DD><<<nT, 128, 0, st>>>(Minv, p, Ap, x, r, z, sc, k, red_partial, red_counter)
      PCG_D(UPD);
      k_pcg_book<<<1, 1, 0, st>>>(sc, k, tol2, max_iter);
      k_pcg_next<<<gv, 256, 0, st>>>(z, p, n, sc, k);
      *launches += 5;
    }
    int h[2] = {0, 0};
    if (cudaMemcpyAsync(h, &sc->done, 2 * sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
    if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
    done = h[0]; iters = h[1];
    if (done) break;
  }
#undef PCG_D
#undef PRE
#undef INI
#undef UPD
  if (iters_out) *iters_out = iters;
  cudaMemcpyAsync(x_out, x, sizeof(double) * n, cudaMemcpyDeviceToDevice, st);
  if (done == 2) return 2;
  // converged (the residual test) or out of iterations: tell them apart by the residual
  double h2[2];
  cudaMemcpyAsync(h2, &sc->rr, 2 * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
  return (h2[0] <= tol2 * h2[1]) ? 0 : 1;
}

}  // namespace bagpu
