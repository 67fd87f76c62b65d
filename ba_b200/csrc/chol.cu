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
// the host).  Factorisation  S = U^T Sigma U,  Sigma = diag(+-1).
//
// Execution: ONE persistent cooperative kernel per solve, dataflow over tile tasks.  A banded S
// (sequential video, configs 2 and 4) is a chain of nT dependent tile rows, so what matters is
// the latency of one link, not FLOPs: a kernel launch per row (the previous version) cost
// ~45 us per row.  Here CTAs claim tasks in a topological order from one counter and spin on
// per-tile ready flags in global memory (release/acquire), so a link costs a flag hop:
//   left-looking tile task (i, j):  R = A_ij - sum_k U_ki^T Sigma_k U_kj  over the rows k that
//       hold both tiles (accumulated in registers while the newest U_ki is still being made);
//     i == j : R = L D L^T by root-free elimination of [R | I | b] in shared memory
//              (one barrier per column) -> U_kk^-1, Sigma, y_k, z_k
//     i <  j : U_ij = Sigma U_ii^-T R (GEMM), c_ij = U_ij^T y_i (its share of the forward
//              substitution, summed in fixed order by the diagonal task of column j)
//   backward substitution, rows bottom-up: tile tasks p_kj = U_kj x_j (tile prefetched into
//       registers before x_j exists), row task x_k = U_kk^-1 (z_k - sum_j p_kj).
// Every tile is written by exactly one task and all sums run in a fixed order: the solve is
// bit-reproducible.  All T x T x T products run on the FP64 tensor pipe (mma.sync.m8n8k4.f64 —
// tcgen05 has no FP64 kind, so DMMA is the FP64 tensor path on sm_100a), fragments loaded
// straight from L2 (ld.global.cg: the tiles are produced by other SMs).
#include <algorithm>
#include <vector>
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"

namespace bagpu {

BA_D void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
BA_D int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
BA_D void st_release(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
BA_D double ldcg(const double* p) { return __ldcg(p); }
// reciprocal off the critical path of the pivot chain: MUFU seed + two Newton steps (~1 ulp)
BA_D double fast_rcp(double d) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
  double e = fma(-d, x, 1.0);
  x = fma(x, e, x);
  e = fma(-d, x, 1.0);
  x = fma(x, e, x);
  return x;
}
BA_D unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define CHOL_STAMP(slot) do { if (c.trace && tid == 0) c.trace[(size_t)t * 4 + (slot)] = gtime(); } while (0)

enum { kTaskDiag = 0, kTaskOff = 1, kTaskBackTile = 2, kTaskBackRow = 3 };

struct CholDev {
  double* tiles;        // S tiles (off-diagonal tiles are overwritten by U_kj)
  double* uinv;         // nT tiles: U_kk^-1 (upper, column-major)
  double* z;            // nT*T: Sigma U^-T b (input of the backward substitution)
  double* sgn;          // nT*T: Sigma
  double* y;            // nT*T: |D|^-1/2 L^-1 b
  double* rhs;          // nT*T: b, then x
  double* cvec;         // n_tiles*T: forward-substitution share of each off-diagonal tile
  double* pvec;         // n_tiles*T: backward-substitution share of each off-diagonal tile
  const CholTask* tasks;
  const CholTerm* terms;
  int* sync;            // [0] task counter, [1] first failing pivot + 1, then tile_flag[n_tiles], diag_flag[nT], x_flag[nT], row_count[nT]
  int n_tasks, n_tiles, nT;
  unsigned long long* trace;   // development aid: 4 globaltimer stamps per task (claim, inputs ready, computed, published)
};

template <int T> struct CholCfg {
  static constexpr int NS = T / 8;                 // 8-wide sub-blocks per tile side
  static constexpr int MH = (NS + 1) / 2;          // sub-block rows per warp item (two halves)
  static constexpr int NT = NS * 2 * 32;           // one warp per (sub-block column, row half)
  static constexpr int LD = T + 1;
  static constexpr int NG2 = NT / T;               // thread groups of the matrix-vector tasks
  static constexpr int XC = (T + NG2 - 1) / NG2;
  static constexpr size_t smem = sizeof(double) * (size_t)(T * LD + 5 * T + NG2 * T);
};

template <int T>
__global__ void __launch_bounds__(CholCfg<T>::NT, 1) k_chol_dataflow(CholDev c) {
  typedef CholCfg<T> Cfg;
  constexpr int LD = Cfg::LD, NS = Cfg::NS, MH = Cfg::MH, NT = Cfg::NT, KQ = T / 4, NG2 = Cfg::NG2, XC = Cfg::XC;
  extern __shared__ double sm[];
  __shared__ int s_task, s_ready;
  double* M = sm;                  // [T][LD]
  double* av = M + T * LD;         // |d|^-1/2
  double* sv = av + T;             // sign(d)
  double* part = sv + T;           // [2][T]
  double* dinv = part + 2 * T;     // 1 / d
  double* part2 = dinv + T;        // [NG2][T]
  int* next = c.sync;
  int* fail = c.sync + 1;
  int* tile_flag = c.sync + 2;
  int* diag_flag = tile_flag + c.n_tiles;
  int* x_flag = diag_flag + c.nT;
  int* row_count = x_flag + c.nT;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
  const int nb = warp % NS, mh = warp / NS;
  const int n0 = nb * 8, mlo = mh * MH, mhi = (mh == 0) ? MH : NS;

  for (;;) {
    if (tid == 0) s_task = atomicAdd(next, 1);
    __syncthreads();
    const int t = s_task;
    __syncthreads();
    if (t >= c.n_tasks) return;
    const CholTask tk = c.tasks[t];
    CHOL_STAMP(0);

    if (tk.kind == kTaskDiag || tk.kind == kTaskOff) {
      // ---- R = A_ij - sum_k U_ki^T Sigma_k U_kj, accumulators in DMMA C-fragment layout ----
      double acc0[MH], acc1[MH];
#pragma unroll
      for (int mi = 0; mi < MH; ++mi) { acc0[mi] = 0.0; acc1[mi] = 0.0; }
      // Terms are listed by ascending row k, the order in which they become ready.  Warp 0 polls the flags of the
      // next (up to 32) terms without blocking and the CTA multiplies out the ready prefix, so that when the newest
      // U_ki arrives everything older is already in the accumulators (waiting for all terms but the newest first
      // left ~60 serial tile products on the critical path of every separator row).
      // diagonal task: the forward-substitution shares c_ik of the same terms are folded in as the terms arrive
      const bool bsum_on = (tk.kind == kTaskDiag) && tid < T;
      double bsum = bsum_on ? c.rhs[(size_t)tk.i * T + tid] : 0.0;
      for (int e = tk.lb; e < tk.le;) {
        if (warp == 0) {
          const int idx = e + lane;
          bool ready = false;
          if (idx < tk.le) {
            const CholTerm tm = c.terms[idx];
            ready = ld_acquire(tile_flag + tm.sa) != 0 && ld_acquire(tile_flag + tm.sb) != 0;
          }
          const unsigned m = __ballot_sync(0xffffffffu, ready);
          if (lane == 0) s_ready = (m == 0xffffffffu) ? 32 : (__ffs(~m) - 1);
        }
        __syncthreads();
        const int nr = s_ready;
        __syncthreads();
        if (nr == 0) { __nanosleep(40); continue; }
        for (int q = 0; q < nr; ++q) {
          const CholTerm tm = c.terms[e + q];
          const double* Ua = c.tiles + (size_t)tm.sa * T * T;
          const double* Ub = c.tiles + (size_t)tm.sb * T * T;
          const double* sg = c.sgn + (size_t)tm.k * T;
          if (bsum_on) bsum -= ldcg(c.cvec + (size_t)tm.sa * T + tid);
          double bfr[KQ];
#pragma unroll
          for (int kq = 0; kq < KQ; ++kq) bfr[kq] = ldcg(Ub + (size_t)(n0 + gid) * T + kq * 4 + tig) * ldcg(sg + kq * 4 + tig);
#pragma unroll
          for (int mi = 0; mi < MH; ++mi) {
            const int mb = mlo + mi;
            if (mb < mhi) {
              double afr[KQ];
#pragma unroll
              for (int kq = 0; kq < KQ; ++kq) afr[kq] = ldcg(Ua + (size_t)(mb * 8 + gid) * T + kq * 4 + tig);
#pragma unroll
              for (int kq = 0; kq < KQ; ++kq) dmma_m8n8k4(acc0[mi], acc1[mi], afr[kq], bfr[kq]);
            }
          }
        }
        e += nr;
      }
      CHOL_STAMP(1);
      double* Aij = c.tiles + (size_t)tk.slot * T * T;
      if (tk.kind == kTaskDiag) {
        const int k = tk.i;
        // R (upper) into M, row-major, zero below the diagonal; column T = b_k - sum c_ik
#pragma unroll
        for (int mi = 0; mi < MH; ++mi) {
          const int mb = mlo + mi;
          if (mb < mhi) {
            const int r = mb * 8 + gid, cc = n0 + 2 * tig;
            M[r * LD + cc] = (r <= cc) ? Aij[(size_t)cc * T + r] - acc0[mi] : 0.0;
            M[r * LD + cc + 1] = (r <= cc + 1) ? Aij[(size_t)(cc + 1) * T + r] - acc1[mi] : 0.0;
          }
        }
        if (bsum_on) M[tid * LD + T] = bsum;
        __syncthreads();
        // Blocked root-free elimination of [R | I | b], pivot blocks of 8, two barriers per block (the
        // column-at-a-time version cost one barrier + one FP64 division per column: 31 us for T = 48).
        //   A  every warp factors the 8x8 diagonal block redundantly with shuffles (lane c owns column c),
        //      so every thread ends up holding the 28 multipliers and 8 pivot reciprocals in registers;
        //   B  one thread per column applies L8^-1 to the 8 block rows (forward substitution in registers);
        //   C  the rows below subtract their 8 multiples of the block rows.
        // Storage as before: upper = transformed rows, strictly lower = L^-1, column T = L^-1 b; pivots in dinv[].
        constexpr int NG = NT / (T + 1), NBK = T / 8;
        const int jg = tid % (T + 1), ig = tid / (T + 1);
        for (int s = 0; s < NBK; ++s) {
          const int p0 = s * 8;
          const int cb = lane & 7;
          double a[8], inv[8], l[28];
#pragma unroll
          for (int r = 0; r < 8; ++r) a[r] = M[(p0 + min(r, cb)) * LD + p0 + max(r, cb)];
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            double d = __shfl_sync(0xffffffffu, a[p], p);
            if (!(d != 0.0) || !isfinite(d)) d = 1.0;
            inv[p] = fast_rcp(d);
#pragma unroll
            for (int i = p + 1; i < 8; ++i) {
              const double li = __shfl_sync(0xffffffffu, a[i], p) * inv[p];
              l[i * (i - 1) / 2 + p] = li;
              a[i] = fma(-li, a[p], a[i]);
            }
          }
          if (warp == 0 && lane < 8) {
            double dd = a[0];
#pragma unroll
            for (int r = 1; r < 8; ++r) if (cb == r) dd = a[r];
            dinv[p0 + cb] = dd;
          }
          // B
          if (tid <= T) {
            const int j = tid;
            const bool ident = (j >= p0 && j < p0 + 8);
            double v[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) v[r] = ident ? ((j - p0 == r) ? 1.0 : 0.0) : M[(p0 + r) * LD + j];
#pragma unroll
            for (int i = 1; i < 8; ++i)
#pragma unroll
              for (int p = 0; p < i; ++p) v[i] = fma(-l[i * (i - 1) / 2 + p], v[p], v[i]);
#pragma unroll
            for (int r = 0; r < 8; ++r)
              if (!ident || r > j - p0) M[(p0 + r) * LD + j] = v[r];
          }
          __syncthreads();
          // C
          if (s < NBK - 1) {
            if (ig < NG) {
              double pj[8];
#pragma unroll
              for (int p = 0; p < 8; ++p) {
                const int jr = jg - p0;
                pj[p] = (jr >= 0 && jr < 8 && jr >= p) ? (jr == p ? 1.0 : 0.0) : M[(p0 + p) * LD + jg];
              }
              for (int i = p0 + 8 + ig; i < T; i += NG) {
                if (jg < p0 + 8 || jg >= i) {
                  double acc = M[i * LD + jg];
#pragma unroll
                  for (int p = 0; p < 8; ++p) acc = fma(-(M[(p0 + p) * LD + i] * inv[p]), pj[p], acc);
                  M[i * LD + jg] = acc;
                }
              }
            }
            __syncthreads();
          }
        }
        for (int r = tid; r < T; r += NT) {
          double d = dinv[r];
          if (!(d != 0.0) || !isfinite(d)) { atomicCAS(fail, 0, k * T + r + 1); d = 1.0; }
          const double a = rsqrt(fabs(d)), s = d < 0.0 ? -1.0 : 1.0;
          av[r] = a; sv[r] = s;
          const double gv = M[r * LD + T];
          c.y[(size_t)k * T + r] = a * gv;
          c.z[(size_t)k * T + r] = s * a * gv;
          c.sgn[(size_t)k * T + r] = s;
        }
        __syncthreads();
        double* xo = c.uinv + (size_t)k * T * T;       // X = U_kk^-1: X[row][col] = a_col * Linv[col][row]
        for (int idx = tid; idx < T * T; idx += NT) {
          const int col = idx / T, row = idx - col * T;
          xo[idx] = (row < col) ? av[col] * M[col * LD + row] : (row == col ? av[col] : 0.0);
        }
        __syncthreads();
        CHOL_STAMP(2);
        if (tid == 0) { __threadfence(); st_release(diag_flag + k, 1); }
        CHOL_STAMP(3);
      } else {
        const int k = tk.i;
        // R into shared memory as the B operand of U_kj = Sigma U_kk^-T R:  Rs[n][kk]
#pragma unroll
        for (int mi = 0; mi < MH; ++mi) {
          const int mb = mlo + mi;
          if (mb < mhi) {
            const int r = mb * 8 + gid, cc = n0 + 2 * tig;
            M[cc * LD + r] = Aij[(size_t)cc * T + r] - acc0[mi];
            M[(cc + 1) * LD + r] = Aij[(size_t)(cc + 1) * T + r] - acc1[mi];
          }
        }
        if (tid == 0) { while (ld_acquire(diag_flag + k) == 0) __nanosleep(40); }
        __syncthreads();
        CHOL_STAMP(2);
        const double* X = c.uinv + (size_t)k * T * T;
        const double* sg = c.sgn + (size_t)k * T;
        const double* yk = c.y + (size_t)k * T;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int mi = 0; mi < MH; ++mi) {
          const int mb = mlo + mi;
          if (mb < mhi) {
            const int r = mb * 8 + gid;
            double c0 = 0.0, c1 = 0.0;
#pragma unroll
            for (int kq = 0; kq < KQ; ++kq) {
              if (kq * 4 <= mb * 8 + 7) {       // U_kk^-T is lower triangular
                const int kk = kq * 4 + tig;
                const double a = (kk <= r) ? ldcg(X + (size_t)r * T + kk) : 0.0;
                dmma_m8n8k4(c0, c1, a, M[(n0 + gid) * LD + kk]);
              }
            }
            const double w = ldcg(sg + r);
            const double u0 = w * c0, u1 = w * c1;
            double* o = Aij + (size_t)(n0 + 2 * tig) * T + r;
            o[0] = u0; o[T] = u1;
            const double yv = ldcg(yk + r);
            s0 += u0 * yv; s1 += u1 * yv;
          }
        }
        // c_kj[col] = sum_r U_kj[r][col] y_k[r]: reduce over the 8 rows of the lane group, then the two halves
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
        if (gid == 0) { part[mh * T + n0 + 2 * tig] = s0; part[mh * T + n0 + 2 * tig + 1] = s1; }
        __syncthreads();
        for (int col = tid; col < T; col += NT) c.cvec[(size_t)tk.slot * T + col] = part[col] + part[T + col];
        __syncthreads();
        if (tid == 0) { __threadfence(); st_release(tile_flag + tk.slot, 1); }
        CHOL_STAMP(3);
      }
    } else {
      // ---- backward substitution -----------------------------------------------------------
      const int g = tid / T, r = tid - g * T;
      const bool on = g < NG2;
      if (tk.kind == kTaskBackTile) {
        const int k = tk.i, j = tk.j;
        if (tid == 0) { while (ld_acquire(tile_flag + tk.slot) == 0) __nanosleep(40); }
        __syncthreads();
        const double* U = c.tiles + (size_t)tk.slot * T * T;
        double ureg[XC];
        if (on) {
#pragma unroll
          for (int q = 0; q < XC; ++q) { const int col = g + q * NG2; ureg[q] = col < T ? ldcg(U + (size_t)col * T + r) : 0.0; }
        }
        if (tid == 0) { while (ld_acquire(x_flag + j) == 0) __nanosleep(40); }
        __syncthreads();
        CHOL_STAMP(1);
        if (on) {
          const double* x = c.rhs + (size_t)j * T;
          double a = 0.0;
#pragma unroll
          for (int q = 0; q < XC; ++q) { const int col = g + q * NG2; if (col < T) a += ureg[q] * ldcg(x + col); }
          part2[g * T + r] = a;
        }
        __syncthreads();
        if (tid < T) {
          double s = 0.0;
#pragma unroll
          for (int gg = 0; gg < NG2; ++gg) s += part2[gg * T + tid];
          c.pvec[(size_t)tk.slot * T + tid] = s;
        }
        __syncthreads();
        if (tid == 0) { __threadfence(); atomicAdd(row_count + k, 1); }
        CHOL_STAMP(3);
      } else {
        const int k = tk.i, m = tk.lb;   // m off-diagonal tiles, slots tk.slot + 1 .. tk.slot + m
        if (tid == 0) { while (ld_acquire(diag_flag + k) == 0) __nanosleep(40); }
        __syncthreads();
        const double* X = c.uinv + (size_t)k * T * T;
        double xreg[XC];
        if (on) {
#pragma unroll
          for (int q = 0; q < XC; ++q) { const int col = g + q * NG2; xreg[q] = (col < T && col >= r) ? ldcg(X + (size_t)col * T + r) : 0.0; }
        }
        if (tid == 0) { while (ld_acquire(row_count + k) < m) __nanosleep(40); __threadfence(); }
        __syncthreads();
        CHOL_STAMP(1);
        if (tid < T) {
          double s = ldcg(c.z + (size_t)k * T + tid);
          for (int e = 1; e <= m; ++e) s -= ldcg(c.pvec + (size_t)(tk.slot + e) * T + tid);
          part[tid] = s;
        }
        __syncthreads();
        if (on) {
          double a = 0.0;
#pragma unroll
          for (int q = 0; q < XC; ++q) { const int col = g + q * NG2; if (col < T) a += xreg[q] * part[col]; }
          part2[g * T + r] = a;
        }
        __syncthreads();
        if (tid < T) {
          double s = 0.0;
#pragma unroll
          for (int gg = 0; gg < NG2; ++gg) s += part2[gg * T + tid];
          c.rhs[(size_t)k * T + tid] = s;
        }
        __syncthreads();
        if (tid == 0) { __threadfence(); st_release(x_flag + k, 1); }
        CHOL_STAMP(3);
      }
    }
  }
}

// ---- host: task list in a topological order (claimed in this order by the CTAs) ---------------
void chol_build_plan(int nT, const std::vector<int>& row_ptr, const std::vector<int>& row_col,
                     std::vector<CholTask>& tasks, std::vector<CholTerm>& terms) {
  tasks.clear(); terms.clear();
  std::vector<std::vector<int>> col_rows(nT);      // rows k < c that hold tile (k, c), ascending
  for (int k = 0; k < nT; ++k)
    for (int e = row_ptr[k]; e < row_ptr[k + 1]; ++e) col_rows[row_col[e]].push_back(k);
  // tiles are numbered row by row, the diagonal tile first, then the row's columns in ascending order
  auto slot = [&](int a, int b) {
    const int base = a + row_ptr[a];
    if (a == b) return base;
    const int* lo = row_col.data() + row_ptr[a];
    const int* hi = row_col.data() + row_ptr[a + 1];
    return base + 1 + (int)(std::lower_bound(lo, hi, b) - lo);
  };
  auto add = [&](int kind, int i, int j, int s, int lb, int le) {
    CholTask t; t.kind = kind; t.i = i; t.j = j; t.slot = s; t.lb = lb; t.le = le; t.pad0 = 0; t.pad1 = 0;
    tasks.push_back(t);
  };
  // Rows are issued wavefront by wavefront of the elimination tree (level = longest dependency chain below the
  // row), so the CTAs' claim window spans every independent branch instead of queueing behind the first one.
  std::vector<int> level(nT, 0), rows(nT);
  for (int i = 0; i < nT; ++i) { for (int k : col_rows[i]) level[i] = std::max(level[i], level[k] + 1); rows[i] = i; }
  std::stable_sort(rows.begin(), rows.end(), [&](int a, int b) { return level[a] < level[b]; });
  for (int ri = 0; ri < nT; ++ri) {
    const int i = rows[ri];
    int lb = (int)terms.size();
    // terms in the order their tiles become ready (wavefront level of the producing row, then row): the kernel
    // accumulates the ready prefix while it waits for the rest
    auto by_level = [&](const CholTerm& a, const CholTerm& b) { return level[a.k] != level[b.k] ? level[a.k] < level[b.k] : a.k < b.k; };
    for (int k : col_rows[i]) { CholTerm tm; tm.sa = slot(k, i); tm.sb = tm.sa; tm.k = k; terms.push_back(tm); }
    std::sort(terms.begin() + lb, terms.end(), by_level);
    add(kTaskDiag, i, i, slot(i, i), lb, (int)terms.size());
    for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e) {
      const int j = row_col[e];
      lb = (int)terms.size();
      const std::vector<int>&ci = col_rows[i], &cj = col_rows[j];
      size_t a = 0, b = 0;
      while (a < ci.size() && b < cj.size()) {
        if (ci[a] == cj[b]) {
          if (ci[a] < i) { CholTerm tm; tm.sa = slot(ci[a], i); tm.sb = slot(ci[a], j); tm.k = ci[a]; terms.push_back(tm); }
          ++a; ++b;
        } else if (ci[a] < cj[b]) ++a; else ++b;
      }
      std::sort(terms.begin() + lb, terms.end(), by_level);
      add(kTaskOff, i, j, slot(i, j), lb, (int)terms.size());
    }
  }
  std::vector<int> level2(nT, 0);
  for (int k = nT - 1; k >= 0; --k) {
    for (int e = row_ptr[k]; e < row_ptr[k + 1]; ++e) level2[k] = std::max(level2[k], level2[row_col[e]] + 1);
    rows[nT - 1 - k] = k;
  }
  std::stable_sort(rows.begin(), rows.end(), [&](int a, int b) { return level2[a] < level2[b]; });
  for (int ri = 0; ri < nT; ++ri) {
    const int k = rows[ri];
    const int m = row_ptr[k + 1] - row_ptr[k];
    for (int e = row_ptr[k]; e < row_ptr[k + 1]; ++e) add(kTaskBackTile, k, row_col[e], slot(k, row_col[e]), 0, 0);
    add(kTaskBackRow, k, k, slot(k, k), m, 0);
  }
}

size_t chol_sync_ints(int n_tiles, int nT) { return 2 + (size_t)n_tiles + 3 * (size_t)nT; }

template <int T>
static int chol_run(CholDev& c, int sms, cudaStream_t st) {
  typedef CholCfg<T> Cfg;
  static int per_sm = 0;
  if (per_sm == 0) {
    cudaFuncSetAttribute(k_chol_dataflow<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_chol_dataflow<T>, Cfg::NT, Cfg::smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  }
  int grid = per_sm * sms;
  if (grid > c.n_tasks) grid = c.n_tasks;
  if (grid < 1) grid = 1;
  void* args[] = {&c};
  return (int)cudaLaunchCooperativeKernel((void*)k_chol_dataflow<T>, dim3(grid), dim3(Cfg::NT), args, Cfg::smem, st);
}

int launch_cholesky_solve(int T, double* tiles, double* uinv, double* work /* 3*nT*T + 2*n_tiles*T */, double* rhs,
                          const CholTask* tasks, const CholTerm* terms, int n_tasks, int* sync_ints, int n_tiles, int nT, int sms,
                          cudaStream_t st, unsigned long long* trace) {
  CholDev c;
  c.trace = trace;
  const size_t v = (size_t)nT * T;
  c.tiles = tiles; c.uinv = uinv; c.z = work; c.sgn = work + v; c.y = work + 2 * v;
  c.cvec = work + 3 * v; c.pvec = c.cvec + (size_t)n_tiles * T; c.rhs = rhs;
  c.tasks = tasks; c.terms = terms; c.sync = sync_ints; c.n_tasks = n_tasks; c.n_tiles = n_tiles; c.nT = nT;
  cudaMemsetAsync(sync_ints, 0, sizeof(int) * chol_sync_ints(n_tiles, nT), st);
  if (T == 48) return chol_run<48>(c, sms, st);
  if (T == 72) return chol_run<72>(c, sms, st);
  if (T == 120) return chol_run<120>(c, sms, st);
  return (int)cudaErrorInvalidValue;
}

}  // namespace bagpu
