// chol.cu — reduced camera system solve  S delta_p = rhs  on the device.
//
// Replaces BundleAdjuster::CalculateGn (reference src/BundleAdjuster.cpp:748-833:
// Eigen::SimplicialLDLT<SparseMatrix, Upper> on s_.sparseView(), or dense LDLT<Upper>).
// S is symmetric positive definite whenever the problem is well posed, so the reference's
// LDL^T and this U^T U factorisation give the same solution up to rounding.
//
// Storage: upper block-tile matrix, tiles of T x T doubles (T = 8 poses x PoseSize), column
// major inside a tile; structurally empty tiles are not stored.  The factorisation is a
// right-looking tiled Cholesky over the tile sparsity (fill computed symbolically on the host):
//   potrf  : U_kk = chol(A_kk), U_kk^-1, and y_k = U_kk^-T b_k       (one CTA)
//   trsm   : U_kj = U_kk^-T A_kj (as a GEMM with the explicit inverse), b_j -= U_kj^T y_k
//   update : A_ij -= U_ki^T U_kj  for every stored pair of row k       (FP64 tensor pipe)
//   back   : x_k = U_kk^-1 (y_k - sum_j U_kj x_j)
// All T x T x T products run on the FP64 tensor pipe (mma.sync.m8n8k4.f64 — tcgen05 has no
// FP64 kind, so DMMA is the FP64 tensor path on sm_100a).  A fully dense S degenerates to the
// dense blocked Cholesky; a banded S (sequential video, configs 2 and 4) skips empty tiles.
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"

namespace bagpu {

constexpr int kKP = 24;       // K-panel depth staged in shared memory
constexpr int kLDP = 28;      // panel leading dimension (28 mod 16 = 12: conflict-free fragment loads)

BA_D void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// C(T x T, column-major in global, ld T) (+)= alpha * A^T B with A, B tiles in global memory
// (column-major, ld T).  The CTA stages K-panels of A and B in shared memory as [col][k].
// mode 0: C = C - A^T B ; mode 1: C = A^T B (overwrite)
template <int T>
BA_D void tile_gemm_atb(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ C,
                        int mode, double* sA, double* sB) {
  constexpr int NS = T / 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int gid = lane >> 2, tig = lane & 3;
  constexpr int MAXSUB = (NS * NS + 7) / 8;   // sub-blocks per warp with >= 8 warps
  double acc0[MAXSUB], acc1[MAXSUB];
#pragma unroll
  for (int s = 0; s < MAXSUB; ++s) { acc0[s] = 0.0; acc1[s] = 0.0; }
  for (int k0 = 0; k0 < T; k0 += kKP) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < T * kKP; idx += blockDim.x) {
      const int col = idx / kKP, kk = idx % kKP;
      sA[col * kLDP + kk] = A[(size_t)col * T + k0 + kk];
      sB[col * kLDP + kk] = B[(size_t)col * T + k0 + kk];
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < MAXSUB; ++s) {
      const int sub = warp + s * nwarps;
      if (sub < NS * NS) {
        const int m0 = (sub % NS) * 8, n0 = (sub / NS) * 8;
        const double* pa = sA + (m0 + gid) * kLDP + tig;
        const double* pb = sB + (n0 + gid) * kLDP + tig;
#pragma unroll
        for (int kk = 0; kk < kKP; kk += 4) dmma_m8n8k4(acc0[s], acc1[s], pa[kk], pb[kk]);
      }
    }
  }
#pragma unroll
  for (int s = 0; s < MAXSUB; ++s) {
    const int sub = warp + s * nwarps;
    if (sub < NS * NS) {
      const int m0 = (sub % NS) * 8, n0 = (sub / NS) * 8;
      double* c = C + (size_t)(n0 + 2 * tig) * T + m0 + gid;
      if (mode == 0) { c[0] -= acc0[s]; c[T] -= acc1[s]; }
      else { c[0] = acc0[s]; c[T] = acc1[s]; }
    }
  }
}

struct CholDev {
  double* tiles;        // S tiles (factored in place)
  double* uinv;         // nT tiles: U_kk^-1 (upper, column-major)
  double* scratch;      // one tile per off-diagonal tile of the widest row (trsm output staging)
  double* rhs;          // nT*T: b -> y -> x
  const int* tile_map;  // nT*nT
  const int* row_ptr;   // nT+1
  const int* row_col;   // off-diagonal tile columns of each row (ascending)
  int* fail;            // first failing tile column + 1
  int nT;
};

// ---- potrf + triangular inverse + forward step of the rhs -------------------------------
template <int T>
__global__ void __launch_bounds__(256) k_potrf(CholDev c, int k) {
  extern __shared__ double sm[];
  constexpr int LD = T + 1;
  // A[col * LD + row]: upper triangle = A_kk -> U_kk, column T = rhs; the strictly lower
  // triangle is reused for U^-1 (X[i][j], i < j, kept at row j / column i), invd = 1 / U_jj.
  double* A = sm;
  double* invd = sm + (T + 1) * LD;
  __shared__ double piv;
  __shared__ int bad;
  double* tile = c.tiles + (size_t)c.tile_map[k * c.nT + k] * T * T;
  for (int idx = threadIdx.x; idx < T * T; idx += blockDim.x) {
    const int col = idx / T, row = idx % T;
    A[col * LD + row] = (row <= col) ? tile[idx] : 0.0;
  }
  for (int r = threadIdx.x; r < T; r += blockDim.x) A[T * LD + r] = c.rhs[(size_t)k * T + r];
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  for (int cidx = 0; cidx < T; ++cidx) {
    if (threadIdx.x == 0) {
      double d = A[cidx * LD + cidx];
      if (!(d > 0.0)) { if (!bad) bad = cidx + 1; d = 1.0; }
      piv = sqrt(d);
      A[cidx * LD + cidx] = piv;
    }
    __syncthreads();
    const double ip = 1.0 / piv;
    for (int j = cidx + 1 + threadIdx.x; j <= T; j += blockDim.x) A[j * LD + cidx] *= ip;   // row cidx of U (and y)
    __syncthreads();
    // trailing update of the upper triangle (and the rhs column): a_ij -= u_ci u_cj, c < i <= j
    const int rem = T - cidx;   // columns cidx+1 .. T
    for (int idx = threadIdx.x; idx < rem * rem; idx += blockDim.x) {
      const int j = cidx + 1 + idx / rem, i = cidx + 1 + idx % rem;
      if (i <= j && i < T) A[j * LD + i] -= A[i * LD + cidx] * A[j * LD + cidx];
    }
    __syncthreads();
  }
  // X = U^-1 (upper): column j by back substitution, one thread per column
  for (int j = threadIdx.x; j < T; j += blockDim.x) {
    const double xjj = 1.0 / A[j * LD + j];
    invd[j] = xjj;
    for (int i = j - 1; i >= 0; --i) {
      double s = A[j * LD + i] * xjj;
      for (int kk = i + 1; kk < j; ++kk) s += A[kk * LD + i] * A[kk * LD + j];
      A[i * LD + j] = -s / A[i * LD + i];
    }
  }
  __syncthreads();
  double* xo = c.uinv + (size_t)k * T * T;
  for (int idx = threadIdx.x; idx < T * T; idx += blockDim.x) {
    const int col = idx / T, row = idx % T;
    tile[idx] = (row <= col) ? A[col * LD + row] : 0.0;
    xo[idx] = (row < col) ? A[row * LD + col] : (row == col ? invd[col] : 0.0);
  }
  for (int r = threadIdx.x; r < T; r += blockDim.x) c.rhs[(size_t)k * T + r] = A[T * LD + r];   // y_k
  if (threadIdx.x == 0 && bad) atomicCAS(c.fail, 0, k * T + bad);
}

// ---- trsm as GEMM: U_kj = (U_kk^-1)^T A_kj ; then b_j -= U_kj^T y_k ------------------------
template <int T>
__global__ void __launch_bounds__(256) k_trsm(CholDev c, int k) {
  extern __shared__ double sm[];
  double* sA = sm;
  double* sB = sm + T * kLDP;
  const int e = c.row_ptr[k] + blockIdx.x;
  const int j = c.row_col[e];
  double* Akj = c.tiles + (size_t)c.tile_map[k * c.nT + j] * T * T;
  double* tmp = c.scratch + (size_t)blockIdx.x * T * T;
  tile_gemm_atb<T>(c.uinv + (size_t)k * T * T, Akj, tmp, 1, sA, sB);
  __syncthreads();
  for (int idx = threadIdx.x; idx < T * T; idx += blockDim.x) Akj[idx] = tmp[idx];
  // rhs: b_j[col] -= sum_r U_kj[r, col] y_k[r]
  const double* y = c.rhs + (size_t)k * T;
  for (int col = threadIdx.x; col < T; col += blockDim.x) {
    double s = 0.0;
    for (int r = 0; r < T; ++r) s += tmp[(size_t)col * T + r] * y[r];
    c.rhs[(size_t)j * T + col] -= s;
  }
}

// ---- trailing update: A_ij -= U_ki^T U_kj for pairs (a <= b) of row k -------------------
template <int T>
__global__ void __launch_bounds__(256) k_update(CholDev c, int k) {
  extern __shared__ double sm[];
  double* sA = sm;
  double* sB = sm + T * kLDP;
  const int base = c.row_ptr[k], m = c.row_ptr[k + 1] - base;
  // decode blockIdx.x -> (a, b), a <= b, enumerated by b
  int pr = blockIdx.x;
  int b = (int)((sqrtf(8.0f * pr + 1.0f) - 1.0f) * 0.5f);
  while ((b + 1) * (b + 2) / 2 <= pr) ++b;
  while (b * (b + 1) / 2 > pr) --b;
  const int a = pr - b * (b + 1) / 2;
  if (b >= m) return;
  const int i = c.row_col[base + a], j = c.row_col[base + b];
  const double* Uki = c.tiles + (size_t)c.tile_map[k * c.nT + i] * T * T;
  const double* Ukj = c.tiles + (size_t)c.tile_map[k * c.nT + j] * T * T;
  double* Aij = c.tiles + (size_t)c.tile_map[i * c.nT + j] * T * T;
  tile_gemm_atb<T>(Uki, Ukj, Aij, 0, sA, sB);
}

// ---- backward substitution, one CTA walks the tile rows bottom-up ---------------------------
template <int T>
__global__ void __launch_bounds__(256) k_backsolve(CholDev c) {
  __shared__ double v[T];
  __shared__ double part[256];
  constexpr int NG = 256 / T > 0 ? 256 / T : 1;   // thread groups splitting the tiles of a row
  const int g = threadIdx.x / T, r = threadIdx.x % T;
  for (int k = c.nT - 1; k >= 0; --k) {
    const int base = c.row_ptr[k], m = c.row_ptr[k + 1] - base;
    double acc = 0.0;
    if (g < NG) {
      for (int e = g; e < m; e += NG) {
        const int j = c.row_col[base + e];
        const double* U = c.tiles + (size_t)c.tile_map[k * c.nT + j] * T * T;
        const double* x = c.rhs + (size_t)j * T;
        for (int col = 0; col < T; ++col) acc += U[(size_t)col * T + r] * x[col];
      }
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x < T) {
      double s = c.rhs[(size_t)k * T + threadIdx.x];
      for (int gg = 0; gg < NG; ++gg) s -= part[gg * T + threadIdx.x];
      v[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x < T) {
      const double* X = c.uinv + (size_t)k * T * T;   // upper: X[col*T + row], row <= col
      double s = 0.0;
      for (int col = threadIdx.x; col < T; ++col) s += X[(size_t)col * T + threadIdx.x] * v[col];
      c.rhs[(size_t)k * T + threadIdx.x] = s;
    }
    __syncthreads();
    __threadfence_block();
  }
}

template <int T>
static void chol_run(const CholDev& c, const int* h_row_ptr, cudaStream_t st, uint64_t* launches) {
  const size_t sm_potrf = sizeof(double) * (size_t)((T + 1) * (T + 1) + T);
  const size_t sm_gemm = sizeof(double) * (size_t)(2 * T * kLDP);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(k_potrf<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_potrf);
    cudaFuncSetAttribute(k_trsm<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_gemm);
    cudaFuncSetAttribute(k_update<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_gemm);
    attr_done = true;
  }
  for (int k = 0; k < c.nT; ++k) {
    k_potrf<T><<<1, 256, sm_potrf, st>>>(c, k);
    ++*launches;
    const int m = h_row_ptr[k + 1] - h_row_ptr[k];
    if (m > 0) {
      k_trsm<T><<<m, 256, sm_gemm, st>>>(c, k);
      k_update<T><<<m * (m + 1) / 2, 256, sm_gemm, st>>>(c, k);
      *launches += 2;
    }
  }
  k_backsolve<T><<<1, 256, 0, st>>>(c);
  ++*launches;
}

void launch_cholesky_solve(int T, double* tiles, double* uinv, double* scratch, double* rhs, const int* tile_map,
                           const int* row_ptr_dev, const int* row_col_dev, const int* row_ptr_host, int nT, int* fail,
                           cudaStream_t st, uint64_t* launches) {
  CholDev c;
  c.tiles = tiles; c.uinv = uinv; c.scratch = scratch; c.rhs = rhs; c.tile_map = tile_map;
  c.row_ptr = row_ptr_dev; c.row_col = row_col_dev; c.fail = fail; c.nT = nT;
  if (T == 48) chol_run<48>(c, row_ptr_host, st, launches);
  else if (T == 72) chol_run<72>(c, row_ptr_host, st, launches);
  else if (T == 120) chol_run<120>(c, row_ptr_host, st, launches);
}

}  // namespace bagpu
