// chol.cu — reduced camera system solve  S delta_p = rhs  on the device.
//
// Replaces BundleAdjuster::CalculateGn (reference src/BundleAdjuster.cpp:748-833:
// Eigen::SimplicialLDLT<SparseMatrix, Upper> on s_.sparseView(), or dense LDLT<Upper>).
// Like the reference's LDL^T the factorisation takes pivots of either sign (a monocular
// problem has a scale gauge direction along which S is numerically singular); only an exactly
// zero / non-finite pivot is a FactorizationError.
//
// Storage: upper block-tile matrix, tiles of T x T doubles (T = 8 poses x PoseSize), column
// major inside a tile; structurally empty tiles are not stored (fill computed symbolically on
// the host).  Factorisation  S = U^T Sigma U,  Sigma = diag(+-1),  right-looking over tile rows:
//   k_factor_row (one CTA per off-diagonal tile of row k, each redoing the small diagonal work):
//       A_kk = L D L^T by root-free elimination of [A_kk | I | b_k] in shared memory
//       (one barrier per column), U_kk = |D|^1/2 L^T, its inverse, z_k = Sigma U_kk^-T b_k;
//       then U_kj = Sigma U_kk^-T A_kj as a GEMM with L^-1 straight from shared memory and
//       b_j -= U_kj^T y_k
//   k_update     : A_ij -= U_ki^T Sigma_k U_kj for every stored pair (i <= j) of row k
//   k_backsolve  : x_k = U_kk^-1 (z_k - sum_j U_kj x_j), one CTA walking the rows bottom-up
// All T x T x T products run on the FP64 tensor pipe (mma.sync.m8n8k4.f64 — tcgen05 has no
// FP64 kind, so DMMA is the FP64 tensor path on sm_100a); fragments are loaded straight from
// global memory / L2 in 32-byte sectors (4 consecutive doubles per lane group), no staging.
// A fully dense S degenerates to the dense blocked factorisation; a banded S (sequential
// video, configs 2 and 4) skips empty tiles and is bound by the chain of tile rows.
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"

namespace bagpu {

BA_D void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

struct CholDev {
  double* tiles;        // S tiles (off-diagonal tiles are overwritten by U_kj)
  double* uinv;         // nT tiles: U_kk^-1 (upper, column-major)
  double* z;            // nT*T: Sigma U^-T b (input of the backward substitution)
  double* sgn;          // nT*T: Sigma
  double* rhs;          // nT*T: b (updated in place), then x
  const int* tile_map;  // nT*nT
  const int* row_ptr;   // nT+1
  const int* row_col;   // off-diagonal tile columns of each row (ascending)
  int* fail;            // first failing row + 1
  int nT;
};

template <int T> struct CholCfg {
  static constexpr int NS = T / 8;                 // 8-wide sub-blocks per tile side
  static constexpr int MH = (NS + 1) / 2;          // sub-block rows per warp item (two halves)
  static constexpr int NT = NS * 2 * 32;           // one warp per (sub-block column, row half)
  static constexpr int LD = T + 1;
  static constexpr size_t smem = sizeof(double) * (size_t)(T * LD + 5 * T);
};

// ---- diagonal tile + one off-diagonal tile of row k --------------------------------------
template <int T>
__global__ void __launch_bounds__(CholCfg<T>::NT, 1) k_factor_row(CholDev c, int k) {
  typedef CholCfg<T> Cfg;
  constexpr int LD = Cfg::LD, NS = Cfg::NS, MH = Cfg::MH, NT = Cfg::NT, KQ = T / 4;
  extern __shared__ double sm[];
  double* M = sm;                  // [T][LD]: upper = A -> D L^T rows, strictly lower = L^-1, column T = rhs
  double* av = M + T * LD;         // |d|^-1/2
  double* sv = av + T;             // sign(d)
  double* part = sv + T;           // [2][T] partial rhs updates
  double* dinv = part + 2 * T;     // 1 / d
  const int tid = threadIdx.x;
  const int m = c.row_ptr[k + 1] - c.row_ptr[k];
  const int j = m ? c.row_col[c.row_ptr[k] + blockIdx.x] : k;
  double* Akj = c.tiles + (size_t)c.tile_map[(size_t)k * c.nT + j] * T * T;
  const int warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
  const int nb = warp % NS, mh = warp / NS;
  const int n0 = nb * 8, mlo = mh * MH, mhi = (mh == 0) ? MH : NS;
  double bfr[KQ];                  // this warp's B fragments of A_kj: in flight during the elimination
#pragma unroll
  for (int kq = 0; kq < KQ; ++kq) bfr[kq] = m ? Akj[(size_t)(n0 + gid) * T + kq * 4 + tig] : 0.0;
  const double* tile = c.tiles + (size_t)c.tile_map[(size_t)k * c.nT + k] * T * T;
  for (int idx = tid; idx < T * T; idx += NT) {
    const int col = idx / T, row = idx - col * T;
    M[row * LD + col] = (row <= col) ? tile[idx] : 0.0;
  }
  for (int r = tid; r < T; r += NT) M[r * LD + T] = c.rhs[(size_t)k * T + r];
  __syncthreads();
  // Root-free elimination of [A | I | b].  The matrix is contiguous (LD = T + 1 columns), thread
  // (ig, jg) owns column jg of the rows cc+1+ig, cc+1+ig+NG, ...: one pivot-row load per step, the
  // multipliers are broadcast loads, and the thread that finishes the next pivot also inverts it,
  // so the FP64 division overlaps the other rows' updates.  One barrier per column.
  constexpr int NG = NT / (T + 1), RPT = (T - 1 + NG - 1) / NG;
  const int jg = tid % (T + 1), ig = tid / (T + 1);
  if (tid == 0) { double d = M[0]; if (!(d != 0.0) || !isfinite(d)) d = 1.0; dinv[0] = 1.0 / d; }
  __syncthreads();
  for (int cc = 0; cc < T - 1; ++cc) {
    const double* rowc = M + cc * LD;
    if (ig < NG) {
      const double inv = dinv[cc];
      const double pj = rowc[jg];
#pragma unroll
      for (int q = 0; q < RPT; ++q) {
        const int i = cc + 1 + ig + q * NG;
        if (i < T) {
          const double f = rowc[i] * inv;
          double* e = M + i * LD + jg;
          if (jg < cc || jg >= i) {
            const double v = *e - f * pj;
            *e = v;
            if (q == 0 && ig == 0 && jg == cc + 1) { double d = v; if (!(d != 0.0) || !isfinite(d)) d = 1.0; dinv[cc + 1] = 1.0 / d; }
          } else if (jg == cc) {
            *e = -f;
          }
        }
      }
    }
    __syncthreads();
  }
  for (int r = tid; r < T; r += NT) {
    double d = M[r * LD + r];
    if (!(d != 0.0) || !isfinite(d)) { atomicCAS(c.fail, 0, k * T + r + 1); d = 1.0; }
    const double a = rsqrt(fabs(d)), s = d < 0.0 ? -1.0 : 1.0;
    av[r] = a; sv[r] = s;
    const double gv = M[r * LD + T];
    M[r * LD + T] = a * gv;                       // y_k
    if (blockIdx.x == 0) { c.z[(size_t)k * T + r] = s * a * gv; c.sgn[(size_t)k * T + r] = s; }
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    double* xo = c.uinv + (size_t)k * T * T;       // X = U_kk^-1: X[row][col] = a_col * Linv[col][row]
    for (int idx = tid; idx < T * T; idx += NT) {
      const int col = idx / T, row = idx - col * T;
      xo[idx] = (row < col) ? av[col] * M[col * LD + row] : (row == col ? av[col] : 0.0);
    }
  }
  if (m == 0) return;
  // ---- U_kj = Sigma |D|^-1/2 L^-1 A_kj, in place (operands were fetched at the top, before any
  // warp of this CTA writes the tile) ---------------------------------------------------------
  double acc0[MH], acc1[MH];
#pragma unroll
  for (int mi = 0; mi < MH; ++mi) { acc0[mi] = 0.0; acc1[mi] = 0.0; }
#pragma unroll
  for (int kq = 0; kq < KQ; ++kq) {
#pragma unroll
    for (int mi = 0; mi < MH; ++mi) {
      const int mb = mlo + mi;
      if (mb < mhi && kq * 4 <= mb * 8 + 7) {       // L^-1 is lower triangular
        const int r = mb * 8 + gid, kk = kq * 4 + tig;
        const double a = (kk < r) ? M[r * LD + kk] : (kk == r ? 1.0 : 0.0);
        dmma_m8n8k4(acc0[mi], acc1[mi], a, bfr[kq]);
      }
    }
  }
  double s0 = 0.0, s1 = 0.0;
#pragma unroll
  for (int mi = 0; mi < MH; ++mi) {
    const int mb = mlo + mi;
    if (mb < mhi) {
      const int r = mb * 8 + gid;
      const double w = sv[r] * av[r];
      const double u0 = w * acc0[mi], u1 = w * acc1[mi];
      double* o = Akj + (size_t)(n0 + 2 * tig) * T + r;
      o[0] = u0; o[T] = u1;
      const double y = M[r * LD + T];
      s0 += u0 * y; s1 += u1 * y;
    }
  }
  // b_j[col] -= sum_r U_kj[r][col] y_k[r]: reduce over the 8 rows of the lane group, then the two halves
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
  if (gid == 0) { part[mh * T + n0 + 2 * tig] = s0; part[mh * T + n0 + 2 * tig + 1] = s1; }
  __syncthreads();
  for (int col = tid; col < T; col += NT) c.rhs[(size_t)j * T + col] -= part[col] + part[T + col];
}

// ---- trailing update: A_ij -= U_ki^T Sigma_k U_kj for pairs (a <= b) of row k ----------------
template <int T>
__global__ void __launch_bounds__(CholCfg<T>::NT, 1) k_update(CholDev c, int k) {
  typedef CholCfg<T> Cfg;
  constexpr int NS = Cfg::NS, MH = Cfg::MH, KQ = T / 4;
  const int base = c.row_ptr[k], m = c.row_ptr[k + 1] - base;
  int pr = blockIdx.x;   // -> (a, b), a <= b, enumerated by b
  int b = (int)((sqrtf(8.0f * pr + 1.0f) - 1.0f) * 0.5f);
  while ((b + 1) * (b + 2) / 2 <= pr) ++b;
  while (b * (b + 1) / 2 > pr) --b;
  const int a = pr - b * (b + 1) / 2;
  if (b >= m) return;
  const int i = c.row_col[base + a], j = c.row_col[base + b];
  const double* Uki = c.tiles + (size_t)c.tile_map[(size_t)k * c.nT + i] * T * T;
  const double* Ukj = c.tiles + (size_t)c.tile_map[(size_t)k * c.nT + j] * T * T;
  double* Aij = c.tiles + (size_t)c.tile_map[(size_t)i * c.nT + j] * T * T;
  const double* sg = c.sgn + (size_t)k * T;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gid = lane >> 2, tig = lane & 3;
  const int nb = warp % NS, mh = warp / NS;
  const int n0 = nb * 8, mlo = mh * MH, mhi = (mh == 0) ? MH : NS;
  double bfr[KQ];
#pragma unroll
  for (int kq = 0; kq < KQ; ++kq) bfr[kq] = Ukj[(size_t)(n0 + gid) * T + kq * 4 + tig] * sg[kq * 4 + tig];
#pragma unroll
  for (int mi = 0; mi < MH; ++mi) {
    const int mb = mlo + mi;
    if (mb < mhi) {
      double afr[KQ];
#pragma unroll
      for (int kq = 0; kq < KQ; ++kq) afr[kq] = Uki[(size_t)(mb * 8 + gid) * T + kq * 4 + tig];
      double c0 = 0.0, c1 = 0.0;
#pragma unroll
      for (int kq = 0; kq < KQ; ++kq) dmma_m8n8k4(c0, c1, afr[kq], bfr[kq]);
      double* o = Aij + (size_t)(n0 + 2 * tig) * T + mb * 8 + gid;
      o[0] -= c0; o[T] -= c1;
    }
  }
}

// ---- backward substitution, one CTA walks the tile rows bottom-up ---------------------------
template <int T>
__global__ void __launch_bounds__(1024, 1) k_backsolve(CholDev c) {
  constexpr int NG = 1024 / T;                     // thread groups, each owning a slice of the columns
  __shared__ double part[NG * T];
  __shared__ double v[T];
  const int g = threadIdx.x / T, r = threadIdx.x - g * T;
  const bool on = g < NG;
  for (int k = c.nT - 1; k >= 0; --k) {
    const int base = c.row_ptr[k], m = c.row_ptr[k + 1] - base;
    // U_kk^-1 does not depend on x: fetch this thread's slice before the dependent part
    constexpr int XC = (T + NG - 1) / NG;
    double xreg[XC];
    const double* X = c.uinv + (size_t)k * T * T;   // upper: X[col*T + row], row <= col
    if (on) {
#pragma unroll
      for (int q = 0; q < XC; ++q) { const int col = g + q * NG; xreg[q] = (col < T && col >= r) ? X[(size_t)col * T + r] : 0.0; }
    }
    double acc = 0.0;
    if (on) {
      for (int e = 0; e < m; ++e) {
        const int j = c.row_col[base + e];
        const double* U = c.tiles + (size_t)c.tile_map[(size_t)k * c.nT + j] * T * T;
        const double* x = c.rhs + (size_t)j * T;
#pragma unroll
        for (int q = 0; q < XC; ++q) { const int col = g + q * NG; if (col < T) acc += U[(size_t)col * T + r] * x[col]; }
      }
      part[g * T + r] = acc;
    }
    __syncthreads();
    if (threadIdx.x < T) {
      double s = c.z[(size_t)k * T + threadIdx.x];
#pragma unroll
      for (int gg = 0; gg < NG; ++gg) s -= part[gg * T + threadIdx.x];
      v[threadIdx.x] = s;
    }
    __syncthreads();
    if (on) {
      double a2 = 0.0;
#pragma unroll
      for (int q = 0; q < XC; ++q) { const int col = g + q * NG; if (col < T) a2 += xreg[q] * v[col]; }
      part[g * T + r] = a2;
    }
    __syncthreads();
    if (threadIdx.x < T) {
      double s = 0.0;
#pragma unroll
      for (int gg = 0; gg < NG; ++gg) s += part[gg * T + threadIdx.x];
      c.rhs[(size_t)k * T + threadIdx.x] = s;
    }
    __syncthreads();
  }
}

template <int T>
static void chol_run(const CholDev& c, const int* h_row_ptr, cudaStream_t st, uint64_t* launches) {
  typedef CholCfg<T> Cfg;
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(k_factor_row<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem);
    attr_done = true;
  }
  for (int k = 0; k < c.nT; ++k) {
    const int m = h_row_ptr[k + 1] - h_row_ptr[k];
    k_factor_row<T><<<m > 0 ? m : 1, Cfg::NT, Cfg::smem, st>>>(c, k);
    ++*launches;
    if (m > 0) {
      k_update<T><<<m * (m + 1) / 2, Cfg::NT, 0, st>>>(c, k);
      ++*launches;
    }
  }
  k_backsolve<T><<<1, 1024, 0, st>>>(c);
  ++*launches;
}

void launch_cholesky_solve(int T, double* tiles, double* uinv, double* work, double* rhs, const int* tile_map,
                           const int* row_ptr_dev, const int* row_col_dev, const int* row_ptr_host, int nT, int* fail,
                           cudaStream_t st, uint64_t* launches) {
  CholDev c;
  c.tiles = tiles; c.uinv = uinv; c.z = work; c.sgn = work + (size_t)nT * T; c.rhs = rhs; c.tile_map = tile_map;
  c.row_ptr = row_ptr_dev; c.row_col = row_col_dev; c.fail = fail; c.nT = nT;
  if (T == 48) chol_run<48>(c, row_ptr_host, st, launches);
  else if (T == 72) chol_run<72>(c, row_ptr_host, st, launches);
  else if (T == 120) chol_run<120>(c, row_ptr_host, st, launches);
}

}  // namespace bagpu
