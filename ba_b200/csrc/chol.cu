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
//   backward substitution: consecutive rows that depend on each other (tile (i, i+1) exists: a dense separator of the
//       nested dissection, a band) form a CHAIN that ONE CTA solves bottom-up without leaving the SM — x of the chain stays
//       in shared memory, the chain's own tiles are streamed from L2 — because a dependent row costs two flag hops (~8 us)
//       when every row is a task of its own and ~2 us inside a chain.  Tiles whose column lies outside the chain are
//       tile tasks p_kj = U_kj x_j as before (prefetched into registers before x_j exists), run by other CTAs.
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
BA_D void cp_async16(void* dst_shared, const void* src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_shared);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}
BA_D void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> BA_D void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
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

enum { kTaskDiag = 0, kTaskOff = 1, kTaskBackTile = 2, kTaskBackRow = 3, kTaskBackChain = 4 };
constexpr int kMaxChain = 16;   // rows per backward chain (their x live in shared memory)
// the 960 threads of T = 120 have 64 registers each: one prefetched tile, not three
constexpr int chol_chain_near(int T) { return T <= 72 ? 3 : 1; }

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
  const int* row_ptr;   // CSR of the off-diagonal tile columns per row (ascending), slot(k, k) = k + row_ptr[k]
  const int* row_col;
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
  // Operand tiles of the T x T x T products are staged in shared memory with cp.async (two stages of two tiles, column
  // stride SS: SS mod 16 = +-4 keeps the DMMA fragment loads of a half-warp on 16 different banks).  The elimination
  // workspace M is only live after the products, so it shares the space.  T = 120 does not fit and loads fragments from L2.
  static constexpr bool STAGED = T <= 72;
  static constexpr int NEAR = chol_chain_near(T);  // columns next to the diagonal that a backward chain multiplies itself
  static constexpr int SS = T + 4;
  static constexpr int TILE_S = T * SS;            // doubles per staged tile
  static constexpr int STAGE = STAGED ? 4 * TILE_S + 2 * T : 0;
  static constexpr int REGION0 = STAGE > T * LD ? STAGE : T * LD;
  static constexpr size_t smem = sizeof(double) * (size_t)(REGION0 + 5 * T + NG2 * T);
};

template <int T>
__global__ void __launch_bounds__(CholCfg<T>::NT, 1) k_chol_dataflow(CholDev c) {
  typedef CholCfg<T> Cfg;
  constexpr int LD = Cfg::LD, NS = Cfg::NS, MH = Cfg::MH, NT = Cfg::NT, KQ = T / 4, NG2 = Cfg::NG2, XC = Cfg::XC;
  constexpr int SS = Cfg::SS, TILE_S = Cfg::TILE_S;
  extern __shared__ __align__(16) double sm[];
  __shared__ int s_task, s_ready;
  __shared__ int s_tsa[32], s_tsb[32], s_tk[32];   // the polled terms (tile slots, producing row)
  double* M = sm;                  // [T][LD]
  double* stg = sm;                // staged operand tiles [2 stages][A, B][T][SS], then Sigma [2][T]  (STAGED only; aliases M)
  double* sgs = sm + 4 * TILE_S;
  double* av = sm + Cfg::REGION0;  // |d|^-1/2
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
      // A_ij in C-fragment layout, fetched now so that the latency is off the path behind the last term
      const double* Aij = c.tiles + (size_t)tk.slot * T * T;
      double a0[MH], a1[MH];
#pragma unroll
      for (int mi = 0; mi < MH; ++mi) {
        const int mb = mlo + mi;
        a0[mi] = 0.0; a1[mi] = 0.0;
        if (mb < mhi) {
          const int r = mb * 8 + gid, cc = n0 + 2 * tig;
          a0[mi] = Aij[(size_t)cc * T + r]; a1[mi] = Aij[(size_t)(cc + 1) * T + r];
        }
      }
      // one term's operand tiles (and Sigma of the producing row) into stage buffer `buf`; q indexes the polled batch
      auto stage_term = [&](int q, int buf) {
        const int sa = s_tsa[q], sb = s_tsb[q];
        double* dA = stg + (size_t)buf * 2 * TILE_S;
        const double* gA = c.tiles + (size_t)sa * T * T;
        for (int ch = tid; ch < T * T / 2; ch += NT) {
          const int col = ch / (T / 2), w = ch - col * (T / 2);
          cp_async16(dA + col * SS + 2 * w, gA + (size_t)col * T + 2 * w);
        }
        if (sb != sa) {
          double* dB = dA + TILE_S;
          const double* gB = c.tiles + (size_t)sb * T * T;
          for (int ch = tid; ch < T * T / 2; ch += NT) {
            const int col = ch / (T / 2), w = ch - col * (T / 2);
            cp_async16(dB + col * SS + 2 * w, gB + (size_t)col * T + 2 * w);
          }
        }
        if (tid < T / 2) cp_async16(sgs + buf * T + 2 * tid, c.sgn + (size_t)s_tk[q] * T + 2 * tid);
        cp_async_commit();
      };
      for (int e = tk.lb; e < tk.le;) {
        if (warp == 0) {
          const int idx = e + lane;
          bool ready = false;
          if (idx < tk.le) {
            const CholTerm tm = c.terms[idx];
            ready = ld_acquire(tile_flag + tm.sa) != 0 && ld_acquire(tile_flag + tm.sb) != 0;
            s_tsa[lane] = tm.sa; s_tsb[lane] = tm.sb; s_tk[lane] = tm.k;
          }
          const unsigned m = __ballot_sync(0xffffffffu, ready);
          if (lane == 0) s_ready = (m == 0xffffffffu) ? 32 : (__ffs(~m) - 1);
        }
        __syncthreads();
        const int nr = s_ready;
        if (nr == 0) { __syncthreads(); __nanosleep(40); continue; }
        if constexpr (Cfg::STAGED) {
          stage_term(0, 0);
          for (int q = 0; q < nr; ++q) {
            const double cv = bsum_on ? ldcg(c.cvec + (size_t)s_tsa[q] * T + tid) : 0.0;
            if (q + 1 < nr) { stage_term(q + 1, (q + 1) & 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
            __syncthreads();
            const double* As = stg + (size_t)(q & 1) * 2 * TILE_S;
            const double* Bs = (s_tsb[q] != s_tsa[q]) ? As + TILE_S : As;
            const double* sg = sgs + (q & 1) * T;
            double bfr[KQ];
#pragma unroll
            for (int kq = 0; kq < KQ; ++kq) bfr[kq] = Bs[(n0 + gid) * SS + kq * 4 + tig] * sg[kq * 4 + tig];
#pragma unroll
            for (int mi = 0; mi < MH; ++mi) {
              const int mb = mlo + mi;
              if (mb < mhi) {
#pragma unroll
                for (int kq = 0; kq < KQ; ++kq) dmma_m8n8k4(acc0[mi], acc1[mi], As[(mb * 8 + gid) * SS + kq * 4 + tig], bfr[kq]);
              }
            }
            bsum -= cv;
            __syncthreads();     // the buffer is refilled two terms on; s_t* by the next poll
          }
        } else {
          for (int q = 0; q < nr; ++q) {
            const int sa = s_tsa[q], sb = s_tsb[q];
            const double* Ua = c.tiles + (size_t)sa * T * T;
            const double* Ub = c.tiles + (size_t)sb * T * T;
            const double* sg = c.sgn + (size_t)s_tk[q] * T;
            if (bsum_on) bsum -= ldcg(c.cvec + (size_t)sa * T + tid);
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
          __syncthreads();
        }
        e += nr;
      }
      CHOL_STAMP(1);
      double* Aout = c.tiles + (size_t)tk.slot * T * T;
      if (tk.kind == kTaskDiag) {
        const int k = tk.i;
        // R (upper) into M, row-major, zero below the diagonal; column T = b_k - sum c_ik
#pragma unroll
        for (int mi = 0; mi < MH; ++mi) {
          const int mb = mlo + mi;
          if (mb < mhi) {
            const int r = mb * 8 + gid, cc = n0 + 2 * tig;
            M[r * LD + cc] = (r <= cc) ? a0[mi] - acc0[mi] : 0.0;
            M[r * LD + cc + 1] = (r <= cc + 1) ? a1[mi] - acc1[mi] : 0.0;
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
            M[cc * LD + r] = a0[mi] - acc0[mi];
            M[(cc + 1) * LD + r] = a1[mi] - acc1[mi];
          }
        }
        if (tid == 0) { while (ld_acquire(diag_flag + k) == 0) __nanosleep(40); }
        __syncthreads();
        CHOL_STAMP(2);
        const double* X = c.uinv + (size_t)k * T * T;
        const double* sg = c.sgn + (size_t)k * T;
        const double* yk = c.y + (size_t)k * T;
        double* Xs = stg + TILE_S;       // behind M
        if constexpr (Cfg::STAGED) {
          for (int ch = tid; ch < T * T / 2; ch += NT) {
            const int col = ch / (T / 2), w = ch - col * (T / 2);
            cp_async16(Xs + col * SS + 2 * w, X + (size_t)col * T + 2 * w);
          }
          cp_async_commit();
        }
        double wv[MH], yv[MH];
#pragma unroll
        for (int mi = 0; mi < MH; ++mi) {
          const int mb = mlo + mi;
          wv[mi] = 0.0; yv[mi] = 0.0;
          if (mb < mhi) { wv[mi] = ldcg(sg + mb * 8 + gid); yv[mi] = ldcg(yk + mb * 8 + gid); }
        }
        if constexpr (Cfg::STAGED) { cp_async_wait<0>(); __syncthreads(); }
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
                double a;
                if constexpr (Cfg::STAGED) a = (kk <= r) ? Xs[r * SS + kk] : 0.0;
                else a = (kk <= r) ? ldcg(X + (size_t)r * T + kk) : 0.0;
                dmma_m8n8k4(c0, c1, a, M[(n0 + gid) * LD + kk]);
              }
            }
            const double u0 = wv[mi] * c0, u1 = wv[mi] * c1;
            double* o = Aout + (size_t)(n0 + 2 * tig) * T + r;
            o[0] = u0; o[T] = u1;
            s0 += u0 * yv[mi]; s1 += u1 * yv[mi];
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
    } else if (tk.kind == kTaskBackChain) {
      // ---- backward substitution of a chain of dependent rows k_hi .. k_lo, bottom-up, on this SM ---------------------------
      // Row i's tiles in the columns i+1 .. i+NEAR of the chain are multiplied here (x from shared memory, the tiles, U_ii^-1 and
      // z_i prefetched into registers one row ahead); every other tile of the row is a kTaskBackTile of another CTA, whose
      // x_j was published at least NEAR rows earlier.
      constexpr int NEAR = Cfg::NEAR;
      const int g = tid / T, r = tid - g * T;
      const bool on = g < NG2;
      const int k_hi = tk.i, k_lo = tk.j, n_rows = k_hi - k_lo + 1;
      double* xs = M;                              // x of the chain's rows: xs[(i - k_lo) * T + c]
      double* part3 = M + kMaxChain * T;           // [NG2][T]
      __shared__ int s_rp[kMaxChain + 1], s_rc[kMaxChain * 3];
      // once diag k_hi is published so is everything the factorisation wrote for the chain's rows (diag i+1 depends on tile (i, i+1))
      if (tid == 0) { while (ld_acquire(diag_flag + k_hi) == 0) __nanosleep(40); }
      if (tid <= n_rows) s_rp[tid] = c.row_ptr[k_lo + tid];
      __syncthreads();
      if (tid < n_rows * NEAR) {
        const int ri = tid / NEAR, q = tid - ri * NEAR;
        const int e = s_rp[ri] + q;
        const int col = e < s_rp[ri + 1] ? c.row_col[e] : 0x7fffffff;
        s_rc[ri * 3 + q] = (col <= k_hi && col <= k_lo + ri + NEAR) ? col : 0x7fffffff;
      }
      __syncthreads();
      double ureg[NEAR][XC], xreg[XC], zreg = 0.0;
      auto load_row = [&](int i, double (&u)[NEAR][XC], double (&x)[XC], double& z) {
        const int ri = i - k_lo, base = i + s_rp[ri];
#pragma unroll
        for (int q = 0; q < NEAR; ++q) {
          const bool have = s_rc[ri * 3 + q] != 0x7fffffff;
          const double* U = c.tiles + (size_t)(base + 1 + q) * T * T;
#pragma unroll
          for (int qq = 0; qq < XC; ++qq) { const int col = g + qq * NG2; u[q][qq] = (on && have && col < T) ? ldcg(U + (size_t)col * T + r) : 0.0; }
        }
        const double* X = c.uinv + (size_t)i * T * T;
#pragma unroll
        for (int qq = 0; qq < XC; ++qq) { const int col = g + qq * NG2; x[qq] = (on && col < T && col >= r) ? ldcg(X + (size_t)col * T + r) : 0.0; }
        z = tid < T ? ldcg(c.z + (size_t)i * T + tid) : 0.0;
      };
      load_row(k_hi, ureg, xreg, zreg);
      for (int i = k_hi; i >= k_lo; --i) {
        const int ri = i - k_lo, base = i + s_rp[ri];
        int n_in = 0;
#pragma unroll
        for (int q = 0; q < NEAR; ++q) n_in += s_rc[ri * 3 + q] != 0x7fffffff;
        const int n_all = s_rp[ri + 1] - s_rp[ri];
        if (on) {
          double a = 0.0;
#pragma unroll
          for (int q = 0; q < NEAR; ++q) {
            const int cj = s_rc[ri * 3 + q];
            if (cj != 0x7fffffff) {
              const double* xj = xs + (size_t)(cj - k_lo) * T;
#pragma unroll
              for (int qq = 0; qq < XC; ++qq) { const int col = g + qq * NG2; if (col < T) a += ureg[q][qq] * xj[col]; }
            }
          }
          part2[g * T + r] = a;
        }
        const double zi = zreg;
        double xnext[XC];
        if (i > k_lo) load_row(i - 1, ureg, xnext, zreg);
        if (tid == 0 && n_all > n_in) { while (ld_acquire(row_count + i) < n_all - n_in) __nanosleep(40); __threadfence(); }
        __syncthreads();
        if (tid < T) {
          double s = zi;
          for (int e = n_in; e < n_all; ++e) s -= ldcg(c.pvec + (size_t)(base + 1 + e) * T + tid);
#pragma unroll
          for (int gg = 0; gg < NG2; ++gg) s -= part2[gg * T + tid];
          part[tid] = s;
        }
        __syncthreads();
        if (on) {
          double b = 0.0;
#pragma unroll
          for (int qq = 0; qq < XC; ++qq) { const int col = g + qq * NG2; if (col < T) b += xreg[qq] * part[col]; }
          part3[g * T + r] = b;
        }
        if (i > k_lo) {
#pragma unroll
          for (int qq = 0; qq < XC; ++qq) xreg[qq] = xnext[qq];
        }
        __syncthreads();
        if (tid < T) {
          double s = 0.0;
#pragma unroll
          for (int gg = 0; gg < NG2; ++gg) s += part3[gg * T + tid];
          c.rhs[(size_t)i * T + tid] = s;
          xs[(size_t)ri * T + tid] = s;
        }
        __syncthreads();
        if (tid == 0) { __threadfence(); st_release(x_flag + i, 1); }
      }
      CHOL_STAMP(3);
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
void chol_build_plan(int T, int nT, const std::vector<int>& row_ptr, const std::vector<int>& row_col,
                     std::vector<CholTask>& tasks, std::vector<CholTerm>& terms) {
  const int near = chol_chain_near(T);
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
  // Inside a wavefront the diagonal tasks come first: an off-diagonal task claimed together with its diagonal task holds a
  // CTA idle for the whole elimination (~11 us), which halves the throughput of wide wavefronts (hundreds of leaves).
  auto by_level = [&](const CholTerm& a, const CholTerm& b) { return level[a.k] != level[b.k] ? level[a.k] < level[b.k] : a.k < b.k; };
  for (int r0 = 0; r0 < nT;) {
    int r1 = r0;
    while (r1 < nT && level[rows[r1]] == level[rows[r0]]) ++r1;
    for (int ri = r0; ri < r1; ++ri) {
      const int i = rows[ri];
      const int lb = (int)terms.size();
      // terms in the order their tiles become ready (wavefront level of the producing row, then row): the kernel
      // accumulates the ready prefix while it waits for the rest
      for (int k : col_rows[i]) { CholTerm tm; tm.sa = slot(k, i); tm.sb = tm.sa; tm.k = k; terms.push_back(tm); }
      std::sort(terms.begin() + lb, terms.end(), by_level);
      add(kTaskDiag, i, i, slot(i, i), lb, (int)terms.size());
    }
    for (int ri = r0; ri < r1; ++ri) {
      const int i = rows[ri];
      for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e) {
        const int j = row_col[e];
        const int lb = (int)terms.size();
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
    r0 = r1;
  }
  // ---- backward substitution: chains of dependent rows (see the header) ----------------------------------------------------
  // chain = maximal run k_lo .. k_hi (at most kMaxChain rows) in which every row holds the tile (i, i + 1)
  std::vector<int> chain_of(nT, -1), chain_lo, chain_hi;
  for (int k = nT - 1; k >= 0;) {
    int lo = k;
    while (lo - 1 >= 0 && k - (lo - 1) < kMaxChain && row_ptr[lo - 1] < row_ptr[lo] && row_col[row_ptr[lo - 1]] == lo) --lo;
    const int id = (int)chain_lo.size();
    chain_lo.push_back(lo); chain_hi.push_back(k);
    for (int i = lo; i <= k; ++i) chain_of[i] = id;
    k = lo - 1;
  }
  // wavefront level of a chain = longest dependency chain of chains above it (columns outside the chain)
  const int n_chains = (int)chain_lo.size();
  std::vector<int> clevel(n_chains, 0), order(n_chains);
  for (int cid = 0; cid < n_chains; ++cid) {       // chains were created from the last row down: their outside columns belong to earlier ids
    for (int i = chain_lo[cid]; i <= chain_hi[cid]; ++i)
      for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e)
        if (row_col[e] > chain_hi[cid]) clevel[cid] = std::max(clevel[cid], clevel[chain_of[row_col[e]]] + 1);
    order[cid] = cid;
  }
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return clevel[a] < clevel[b]; });
  for (int oi = 0; oi < n_chains; ++oi) {
    const int cid = order[oi], lo = chain_lo[cid], hi = chain_hi[cid];
    // tiles outside the chain first (their x belong to chains listed earlier), then the chain, then the chain's own tiles
    // beyond the near columns: those wait for x the chain publishes as it descends, so they are listed top row first and
    // behind the chain task — every task before them is then either done or held by a CTA that does not wait for them
    for (int i = hi; i >= lo; --i)
      for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e)
        if (row_col[e] > hi) add(kTaskBackTile, i, row_col[e], slot(i, row_col[e]), 0, 0);
    add(kTaskBackChain, hi, lo, slot(hi, hi), 0, 0);
    for (int i = hi; i >= lo; --i)
      for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e)
        if (row_col[e] <= hi && row_col[e] > i + near) add(kTaskBackTile, i, row_col[e], slot(i, row_col[e]), 0, 0);
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
                          const CholTask* tasks, const CholTerm* terms, const int* row_ptr, const int* row_col, int n_tasks, int* sync_ints,
                          int n_tiles, int nT, int sms, cudaStream_t st, unsigned long long* trace) {
  CholDev c;
  c.trace = trace;
  const size_t v = (size_t)nT * T;
  c.tiles = tiles; c.uinv = uinv; c.z = work; c.sgn = work + v; c.y = work + 2 * v;
  c.cvec = work + 3 * v; c.pvec = c.cvec + (size_t)n_tiles * T; c.rhs = rhs;
  c.tasks = tasks; c.terms = terms; c.row_ptr = row_ptr; c.row_col = row_col; c.sync = sync_ints; c.n_tasks = n_tasks; c.n_tiles = n_tiles; c.nT = nT;
  cudaMemsetAsync(sync_ints, 0, sizeof(int) * chol_sync_ints(n_tiles, nT), st);
  if (T == 48) return chol_run<48>(c, sms, st);
  if (T == 72) return chol_run<72>(c, sms, st);
  if (T == 120) return chol_run<120>(c, sms, st);
  return (int)cudaErrorInvalidValue;
}

}  // namespace bagpu
