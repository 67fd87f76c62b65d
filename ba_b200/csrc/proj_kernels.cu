// proj_kernels.cu — kernels and launchers for the projection part of the hot path.
// See proj_kernels.cuh for the reference lines each kernel replaces.
#include "proj_kernels.cuh"
#include "launch.h"

namespace bagpu {

// ---------------------------------------------------------------------------------------
// pose frames: T_sw / T_ws per (pose, camera) and R_wp  (replaces PoseT::GetTsw, Types.h:61-70)
__global__ void k_pose_frames(DeviceProblem P) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P.n_poses * P.n_cams) return;
  const int p = idx / P.n_cams, c = idx % P.n_cams;
  double Rwp[9];
  quat_to_R(P.pose_q + 4 * (size_t)p, Rwp);
  const double* t = P.pose_t + 3 * (size_t)p;
  if (c == 0) {
#pragma unroll
    for (int i = 0; i < 9; ++i) P.Rwp[9 * (size_t)p + i] = Rwp[i];
  }
  double Rws[9], tws[3];
  mat3_mul(Rwp, P.cam_Rvs + 9 * c, Rws);
  mat3_vec(Rwp, P.cam_tvs + 3 * c, tws);
  tws[0] += t[0]; tws[1] += t[1]; tws[2] += t[2];
  double* o = P.Tws + (size_t)idx * 12;
#pragma unroll
  for (int i = 0; i < 9; ++i) o[i] = Rws[i];
  o[9] = tws[0]; o[10] = tws[1]; o[11] = tws[2];
  double* s = P.Tsw + (size_t)idx * 12;
  // R_sw = R_ws^T, t_sw = -R_sw t_ws
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) s[3 * r + cc] = Rws[3 * cc + r];
  double tsw[3];
  mat3T_vec(Rws, tws, tsw);
  s[9] = -tsw[0]; s[10] = -tsw[1]; s[11] = -tsw[2];
}

// Solve() prologue for LmSize 1: x_s = T_sw(ref) x_w, normalised by |x_s.xyz|  (cpp:287-296)
__global__ void k_prepare_xs(DeviceProblem P, const double* lm_xw) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  const double* xw = lm_xw + 4 * (size_t)lm;
  const double* Tsw = P.Tsw + ((size_t)P.lm_ref_pose[lm] * P.n_cams + P.lm_ref_cam[lm]) * 12;
  double xs[3];
  mat3_vec(Tsw, xw, xs);
  xs[0] += Tsw[9] * xw[3]; xs[1] += Tsw[10] * xw[3]; xs[2] += Tsw[11] * xw[3];
  const double len = sqrt(xs[0] * xs[0] + xs[1] * xs[1] + xs[2] * xs[2]);
  double* o = P.lm_x + 4 * (size_t)lm;
  o[0] = xs[0] / len; o[1] = xs[1] / len; o[2] = xs[2] / len; o[3] = xw[3] / len;
}

// Solve() epilogue for LmSize 1: x_w = T_sw(ref)^-1 x_s  (cpp:671-678)
__global__ void k_finish_xw(DeviceProblem P, double* lm_xw) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  const double* xs = P.lm_x + 4 * (size_t)lm;
  const double* Tws = P.Tws + ((size_t)P.lm_ref_pose[lm] * P.n_cams + P.lm_ref_cam[lm]) * 12;
  double xw[3];
  mat3_vec(Tws, xs, xw);
  double* o = lm_xw + 4 * (size_t)lm;
  o[0] = xw[0] + Tws[9] * xs[3]; o[1] = xw[1] + Tws[10] * xs[3]; o[2] = xw[2] + Tws[11] * xs[3]; o[3] = xs[3];
}

// ---------------------------------------------------------------------------------------
// pass 1: residual -> maha0 = |r|^2 * orig_weight   (parallel_algos.h:143-150)
template <int L, int G>
__global__ void __launch_bounds__(256) k_residual(DeviceProblem P) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int group = (blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int n_groups = (gridDim.x * blockDim.x) / G;
  for (int lm = group; lm < P.n_lms; lm += n_groups) {
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    if (s == e) continue;
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    for (int i = s + gl; i < e; i += G) {
      double r[2];
      obs_residual<L>(P, c, i, r);
      P.obs_maha0[i] = (r[0] * r[0] + r[1] * r[1]) * P.obs_w0[i];
    }
  }
}

// ---------------------------------------------------------------------------------------
// pass 2: Jacobians, Huber weights, V / b_l / V^-1, W, U and b_p
template <int L, int G>
__global__ void __launch_bounds__(128) k_accumulate(DeviceProblem P) {
  constexpr int NV = L * (L + 1) / 2;
  const int lane = threadIdx.x & 31, gl = lane % G;
  const unsigned gmask = group_mask<G>(lane);
  const int group = (blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int n_groups = (gridDim.x * blockDim.x) / G;
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median);
  double err = 0.0;
  for (int lm = group; lm < P.n_lms; lm += n_groups) {
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    const bool lm_on = c.lm_opt >= 0;
    double V[NV], bl[L];
#pragma unroll
    for (int i = 0; i < NV; ++i) V[i] = 0.0;
#pragma unroll
    for (int i = 0; i < L; ++i) bl[i] = 0.0;
    // reference-pose accumulators (LmSize 1): H_rr upper (21), g_r (6), W_r (6)
    double Hrr[L == 1 ? 21 : 1], gr[L == 1 ? 6 : 1], Wr[L == 1 ? 6 : 1];
    if (L == 1) {
#pragma unroll
      for (int i = 0; i < 21; ++i) Hrr[i] = 0.0;
#pragma unroll
      for (int i = 0; i < 6; ++i) { gr[i] = 0.0; Wr[i] = 0.0; }
    }
    bool any_ref = false;
    for (int i = s + gl; i < e; i += G) {
      ObsJac<L> o;
      obs_jacobians<L>(P, c, i, o);
      const double w0 = P.obs_w0[i];
      const double maha0 = (o.r[0] * o.r[0] + o.r[1] * o.r[1]) * w0;
      const double w = huber_weight(maha0, w0, c_huber, P.use_robust);
      P.obs_w[i] = w;
      err += (o.r[0] * o.r[0] + o.r[1] * o.r[1]) * w;
      const double wr0 = w * o.r[0], wr1 = w * o.r[1];
      if (lm_on) {
        int k = 0;
#pragma unroll
        for (int a = 0; a < L; ++a) {
#pragma unroll
          for (int b = a; b < L; ++b) { V[k] += w * (o.Jl[a] * o.Jl[b] + o.Jl[L + a] * o.Jl[L + b]); ++k; }
          bl[a] += o.Jl[a] * wr0 + o.Jl[L + a] * wr1;
        }
      }
      double* Wout = P.obs_W + (size_t)i * 6 * L;
      if (o.pm >= 0) {
        // W_m = w J_m^T J_l  (6 x L, row-major)
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
          for (int b = 0; b < L; ++b)
            Wout[a * L + b] = lm_on ? w * (o.Jm[a] * o.Jl[b] + o.Jm[6 + a] * o.Jl[L + b]) : 0.0;
        // U(m,m) += w J_m^T J_m (upper triangle of the diagonal block), b_p[m] += w J_m^T r
        double* blk = s_block(P, o.pm, o.pm);
#pragma unroll
        for (int b = 0; b < 6; ++b)
#pragma unroll
          for (int a = 0; a <= b; ++a) red_add(blk + (size_t)b * P.T + a, w * (o.Jm[a] * o.Jm[b] + o.Jm[6 + a] * o.Jm[6 + b]));
        double* bp = P.b_p + (size_t)o.pm * P.D;
#pragma unroll
        for (int a = 0; a < 6; ++a) red_add(bp + a, o.Jm[a] * wr0 + o.Jm[6 + a] * wr1);
      } else {
#pragma unroll
        for (int a = 0; a < 6 * L; ++a) Wout[a] = 0.0;
      }
      if (L == 1 && o.ref_on) {
        any_ref = true;
        int k = 0;
#pragma unroll
        for (int b = 0; b < 6; ++b) {
#pragma unroll
          for (int a = 0; a <= b; ++a) { Hrr[k] += w * (o.Jr[a] * o.Jr[b] + o.Jr[6 + a] * o.Jr[6 + b]); ++k; }
          gr[b] += o.Jr[b] * wr0 + o.Jr[6 + b] * wr1;
          if (lm_on) Wr[b] += w * (o.Jr[b] * o.Jl[0] + o.Jr[6 + b] * o.Jl[1]);
        }
        if (o.pm >= 0) {
          // cross block of this residual: U(min,max) += w J_min^T J_max  (SparseBlockMatrixOps.h:236-241)
          const bool m_first = o.pm < c.ref_opt;
          double* blk = m_first ? s_block(P, o.pm, c.ref_opt) : s_block(P, c.ref_opt, o.pm);
          const double* Ja = m_first ? o.Jm : o.Jr;
          const double* Jb = m_first ? o.Jr : o.Jm;
#pragma unroll
          for (int b = 0; b < 6; ++b)
#pragma unroll
            for (int a = 0; a < 6; ++a) red_add(blk + (size_t)b * P.T + a, w * (Ja[a] * Jb[b] + Ja[6 + a] * Jb[6 + b]));
        }
      }
    }
    // ---- group reductions -------------------------------------------------------------
#pragma unroll
    for (int i = 0; i < NV; ++i) V[i] = group_sum<G>(V[i], gmask);
#pragma unroll
    for (int i = 0; i < L; ++i) bl[i] = group_sum<G>(bl[i], gmask);
    if (L == 1) {
      const bool grp_ref = __any_sync(gmask, any_ref);
      if (grp_ref) {
#pragma unroll
        for (int i = 0; i < 21; ++i) Hrr[i] = group_sum<G>(Hrr[i], gmask);
#pragma unroll
        for (int i = 0; i < 6; ++i) { gr[i] = group_sum<G>(gr[i], gmask); Wr[i] = group_sum<G>(Wr[i], gmask); }
        // scatter the reference-pose blocks, elements spread over the lanes of the group
        double* blk = s_block(P, c.ref_opt, c.ref_opt);
        int k = 0;
#pragma unroll
        for (int b = 0; b < 6; ++b)
#pragma unroll
          for (int a = 0; a <= b; ++a) { if ((k % G) == gl) red_add(blk + (size_t)b * P.T + a, Hrr[k]); ++k; }
#pragma unroll
        for (int a = 0; a < 6; ++a) if ((a % G) == gl) red_add(P.b_p + (size_t)c.ref_opt * P.D + a, gr[a]);
      }
      if (gl == 0) {
#pragma unroll
        for (int a = 0; a < 6; ++a) P.lm_Wref[(size_t)lm * 6 + a] = (grp_ref && lm_on) ? Wr[a] : 0.0;
      }
    }
    if (gl == 0) {
      double* vo = P.lm_Vinv + (size_t)lm * L * L;
      double* bo = P.lm_bl + (size_t)lm * L;
      if (lm_on) {
        if (L == 1) {
          double v = V[0];
          if (fabs(v) < 1e-6) v += 1e-6;            // cpp:431-434
          vo[0] = 1.0 / v;
          bo[0] = bl[0];
        } else {
          double M[9] = {V[0], V[1], V[2], V[1], V[3], V[4], V[2], V[4], V[5]};
          double n2 = 0.0;
#pragma unroll
          for (int i = 0; i < 9; ++i) n2 += M[i] * M[i];
          if (sqrt(n2) < 1e-6) { M[0] += 1e-6; M[4] += 1e-6; M[8] += 1e-6; }   // cpp:436-439
          double Mi[9];
          inv3(M, Mi);
#pragma unroll
          for (int i = 0; i < 9; ++i) vo[i] = Mi[i];
#pragma unroll
          for (int i = 0; i < L; ++i) bo[i] = bl[i];
        }
      } else {
#pragma unroll
        for (int i = 0; i < L * L; ++i) vo[i] = 0.0;
#pragma unroll
        for (int i = 0; i < L; ++i) bo[i] = 0.0;
      }
    }
  }
  err = warp_sum(err);
  if (lane == 0) red_add(&P.sc->proj_error, err);
}

// ---------------------------------------------------------------------------------------
// Schur-product kernel: S -= W V^-1 W^T, rhs_sc -= W V^-1 b_l, one warp per landmark.
// Entries of a landmark = its observations (+ the reference pose for LmSize 1); they are staged
// NE at a time in shared memory (W and E = W V^-1), and the (pair, element) work items are
// spread over the 32 lanes so that consecutive lanes hit consecutive addresses of one block.
template <int L>
struct SchurStage {
  double W[16][6 * L];
  double E[16][6 * L];
  int P[16];
};

template <int L>
BA_D void schur_stage_load(const DeviceProblem& P, SchurStage<L>& st, int lm, int s, int n_obs_lm, int n_ent,
                           int base, const double* Vinv, int ref_pose, int ref_opt, int lane) {
  constexpr int NE = 16;
  const int cnt = min(NE, n_ent - base);
  for (int t = lane; t < cnt * 6 * L; t += 32) {
    const int e = base + t / (6 * L), idx = t % (6 * L);
    st.W[t / (6 * L)][idx] = (e < n_obs_lm) ? P.obs_W[(size_t)(s + e) * 6 * L + idx] : P.lm_Wref[(size_t)lm * 6 * L + idx];
  }
  if (lane < cnt) {
    const int e = base + lane;
    int po;
    if (e < n_obs_lm) {
      const int m = (int)P.obs_pose[s + e];
      const int mo = P.pose_opt[m];
      po = (m != ref_pose && mo >= 0) ? mo : -1;
    } else {
      po = ref_opt;
    }
    st.P[lane] = po;
  }
  __syncwarp();
  for (int t = lane; t < cnt * 6 * L; t += 32) {
    const int e = t / (6 * L), idx = t % (6 * L), r = idx / L, cc = idx % L;
    double v = 0.0;
#pragma unroll
    for (int k = 0; k < L; ++k) v += st.W[e][r * L + k] * Vinv[k * L + cc];
    st.E[e][idx] = v;
  }
  __syncwarp();
}

template <int L>
__global__ void __launch_bounds__(128) k_schur(DeviceProblem P) {
  constexpr int NE = 16;
  __shared__ SchurStage<L> stA[4], stB[4];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int n_warps = (gridDim.x * blockDim.x) >> 5;
  SchurStage<L>& A = stA[wib];
  for (int lm = warp; lm < P.n_lms; lm += n_warps) {
    if (P.lm_opt[lm] < 0) continue;
    const int s = P.lm_obs_start[lm], n_obs_lm = P.lm_obs_start[lm + 1] - s;
    const int ref_pose = (int)P.lm_ref_pose[lm];
    const int ref_opt = (L == 1) ? P.pose_opt[ref_pose] : -1;
    const int n_ent = n_obs_lm + ((L == 1 && ref_opt >= 0) ? 1 : 0);
    if (n_ent == 0) continue;
    double Vinv[L * L], bl[L];
#pragma unroll
    for (int i = 0; i < L * L; ++i) Vinv[i] = P.lm_Vinv[(size_t)lm * L * L + i];
#pragma unroll
    for (int i = 0; i < L; ++i) bl[i] = P.lm_bl[(size_t)lm * L + i];
    for (int ab = 0; ab < n_ent; ab += NE) {
      schur_stage_load<L>(P, A, lm, s, n_obs_lm, n_ent, ab, Vinv, ref_pose, ref_opt, lane);
      const int na = min(NE, n_ent - ab);
      // rhs_sc[p_a] -= E_a b_l
      for (int t = lane; t < na * 6; t += 32) {
        const int e = t / 6, r = t % 6;
        const int po = A.P[e];
        if (po >= 0) {
          double v = 0.0;
#pragma unroll
          for (int k = 0; k < L; ++k) v += A.E[e][r * L + k] * bl[k];
          red_add(P.rhs_sc + (size_t)po * P.D + r, -v);
        }
      }
      for (int bb = ab; bb < n_ent; bb += NE) {
        SchurStage<L>& B = (bb == ab) ? A : stB[wib];
        if (bb != ab) schur_stage_load<L>(P, B, lm, s, n_obs_lm, n_ent, bb, Vinv, ref_pose, ref_opt, lane);
        const int nb = min(NE, n_ent - bb);
        const int n_pairs = (bb == ab) ? na * (na + 1) / 2 : na * nb;
        for (int t = lane; t < n_pairs * 36; t += 32) {
          const int pr = t / 36, el = t % 36;
          int a, b;
          if (bb == ab) {  // upper-triangular enumeration by column b: pr = b(b+1)/2 + a
            b = (int)((sqrtf(8.0f * pr + 1.0f) - 1.0f) * 0.5f);
            while ((b + 1) * (b + 2) / 2 <= pr) ++b;
            while (b * (b + 1) / 2 > pr) --b;
            a = pr - b * (b + 1) / 2;
          } else {
            b = pr / na; a = pr % na;
          }
          const int pa = A.P[a], pb = B.P[b];
          if (pa < 0 || pb < 0) continue;
          const int r = el % 6, cidx = el / 6;
          double v = 0.0;
          int bi, bj;
          if (pa < pb || (pa == pb && bb == ab && a == b)) {
            bi = pa; bj = pb;
            if (pa == pb && r > cidx) continue;  // diagonal block: upper triangle only
#pragma unroll
            for (int k = 0; k < L; ++k) v += A.E[a][r * L + k] * B.W[b][cidx * L + k];
          } else if (pa > pb) {
            bi = pb; bj = pa;
#pragma unroll
            for (int k = 0; k < L; ++k) v += B.E[b][r * L + k] * A.W[a][cidx * L + k];
          } else {  // same pose seen twice (multi-camera rig): both cross terms land in the diagonal block
            bi = pa; bj = pa;
            if (r > cidx) continue;
#pragma unroll
            for (int k = 0; k < L; ++k) v += A.E[a][r * L + k] * B.W[b][cidx * L + k] + B.E[b][r * L + k] * A.W[a][cidx * L + k];
          }
          red_add(s_block(P, bi, bj) + (size_t)cidx * P.T + r, -v);
        }
        __syncwarp();
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// steepest-descent denominator |J_p b_p + J_l b_l|^2 over projection residuals (cpp:879-906)
template <int L, int G>
__global__ void __launch_bounds__(128) k_sd_denominator(DeviceProblem P) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int group = (blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int n_groups = (gridDim.x * blockDim.x) / G;
  double acc = 0.0;
  for (int lm = group; lm < P.n_lms; lm += n_groups) {
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    if (s == e) continue;
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    double bl[L];
#pragma unroll
    for (int i = 0; i < L; ++i) bl[i] = (c.lm_opt >= 0) ? P.lm_bl[(size_t)lm * L + i] : 0.0;
    for (int i = s + gl; i < e; i += G) {
      ObsJac<L> o;
      obs_jacobians<L>(P, c, i, o);
      double v0 = 0.0, v1 = 0.0;
#pragma unroll
      for (int k = 0; k < L; ++k) { v0 += o.Jl[k] * bl[k]; v1 += o.Jl[L + k] * bl[k]; }
      if (o.pm >= 0) {
        const double* bp = P.b_p + (size_t)o.pm * P.D;
#pragma unroll
        for (int k = 0; k < 6; ++k) { v0 += o.Jm[k] * bp[k]; v1 += o.Jm[6 + k] * bp[k]; }
      }
      if (L == 1 && o.ref_on) {
        const double* bp = P.b_p + (size_t)c.ref_opt * P.D;
#pragma unroll
        for (int k = 0; k < 6; ++k) { v0 += o.Jr[k] * bp[k]; v1 += o.Jr[6 + k] * bp[k]; }
      }
      acc += P.obs_w[i] * (v0 * v0 + v1 * v1);
    }
  }
  acc = warp_sum(acc);
  if (lane == 0) red_add(&P.sc->sd_den_proj, acc);
}

// ---------------------------------------------------------------------------------------
// back-substitution: delta_l = V^-1 (b_l - W^T delta_p)   (cpp:709-744)
template <int L, int G>
__global__ void __launch_bounds__(256) k_back_substitute(DeviceProblem P, const double* delta_p, double* delta_l) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const unsigned gmask = group_mask<G>(lane);
  const int group = (blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int n_groups = (gridDim.x * blockDim.x) / G;
  for (int lm = group; lm < P.n_lms; lm += n_groups) {
    const int lo = P.lm_opt[lm];
    if (lo < 0) continue;
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    const int ref_pose = (int)P.lm_ref_pose[lm];
    double acc[L];
#pragma unroll
    for (int k = 0; k < L; ++k) acc[k] = 0.0;
    for (int i = s + gl; i < e; i += G) {
      const int m = (int)P.obs_pose[i];
      const int mo = P.pose_opt[m];
      if (m == ref_pose || mo < 0) continue;
      const double* W = P.obs_W + (size_t)i * 6 * L;
      const double* dp = delta_p + (size_t)mo * P.D;
#pragma unroll
      for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int k = 0; k < L; ++k) acc[k] += W[a * L + k] * dp[a];
    }
#pragma unroll
    for (int k = 0; k < L; ++k) acc[k] = group_sum<G>(acc[k], gmask);
    if (gl == 0) {
      if (L == 1) {
        const int ro = P.pose_opt[ref_pose];
        if (ro >= 0) {
          const double* W = P.lm_Wref + (size_t)lm * 6;
          const double* dp = delta_p + (size_t)ro * P.D;
#pragma unroll
          for (int a = 0; a < 6; ++a) acc[0] += W[a] * dp[a];
        }
      }
      double rhs[L];
#pragma unroll
      for (int k = 0; k < L; ++k) rhs[k] = P.lm_bl[(size_t)lm * L + k] - acc[k];
      const double* Vi = P.lm_Vinv + (size_t)lm * L * L;
#pragma unroll
      for (int r = 0; r < L; ++r) {
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < L; ++k) v += Vi[r * L + k] * rhs[k];
        delta_l[(size_t)lo * L + r] = v;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// residual-only evaluation with the current Huber weights (cpp:148-188)
template <int L, int G>
__global__ void __launch_bounds__(256) k_evaluate(DeviceProblem P, double* out_sum, double* res_out, double* maha_out) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int group = (blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int n_groups = (gridDim.x * blockDim.x) / G;
  double acc = 0.0;
  const double thr2 = P.outlier_threshold * P.outlier_threshold;
  for (int lm = group; lm < P.n_lms; lm += n_groups) {
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    if (s == e) continue;
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    unsigned n_out = 0;
    for (int i = s + gl; i < e; i += G) {
      double r[2];
      obs_residual<L>(P, c, i, r);
      const double n2 = r[0] * r[0] + r[1] * r[1];
      const double m = n2 * P.obs_w[i];
      acc += m;
      if (sqrt(n2) > P.outlier_threshold) ++n_out;
      if (res_out) {
        const uint32_t id = P.obs_perm[i];
        res_out[2 * (size_t)id] = r[0]; res_out[2 * (size_t)id + 1] = r[1];
        maha_out[id] = m;
      }
    }
    if (n_out) atomicAdd(&P.lm_outliers[lm], n_out);
  }
  (void)thr2;
  acc = warp_sum(acc);
  if (lane == 0) red_add(out_sum, acc);
}

// ---------------------------------------------------------------------------------------
// ApplyUpdate (cpp:84-139): poses and landmarks, coefficient `coef`
__global__ void k_update_poses(DeviceProblem P, const double* delta_p, double coef) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P.n_poses) return;
  const int o = P.pose_opt[p];
  if (o < 0) return;
  const double* d = delta_p + (size_t)o * P.D;
  double* t = P.pose_t + 3 * (size_t)p;
  double* q = P.pose_q + 4 * (size_t)p;
  t[0] += -d[0] * coef; t[1] += -d[1] * coef; t[2] += -d[2] * coef;
  const double w[3] = {-d[3] * coef, -d[4] * coef, -d[5] * coef};
  double dq[4], qn[4];
  so3_exp(w, dq);
  quat_mul(q, dq, qn);
  quat_normalize(qn);
  q[0] = qn[0]; q[1] = qn[1]; q[2] = qn[2]; q[3] = qn[3];
  if (P.D >= 9) { double* v = P.pose_v + 3 * (size_t)p; v[0] -= d[6] * coef; v[1] -= d[7] * coef; v[2] -= d[8] * coef; }
  if (P.D >= 15) { double* b = P.pose_b + 6 * (size_t)p; for (int k = 0; k < 6; ++k) b[k] -= d[9 + k] * coef; }
}

template <int L>
__global__ void k_update_landmarks(DeviceProblem P, const double* delta_l, double coef) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  const int o = P.lm_opt[lm];
  if (o < 0) return;
  double* x = P.lm_x + 4 * (size_t)lm;
  if (L == 1) {
    const double d = delta_l[o] * coef;
    x[3] -= d;
    if (x[3] < 0) { x[3] += d; P.lm_reliable[lm] = 0; }   // cpp:127-134
  } else {
#pragma unroll
    for (int k = 0; k < L; ++k) x[k] -= delta_l[(size_t)o * L + k] * coef;
  }
}

// masked parameters: S(idx, idx) = 1e6 (cpp:586-598); padding rows of the last tile: 1
__global__ void k_mask_diagonal(DeviceProblem P, const uint8_t* mask_used, int global_flag) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_pad = P.nT * P.T;
  if (idx >= n_pad) return;
  const int n = P.n_active_poses * P.D;
  const int tI = idx / P.T;
  double* diag = P.S_tiles + (size_t)P.tile_map[tI * P.nT + tI] * P.T * P.T + (size_t)(idx % P.T) * P.T + (idx % P.T);
  if (idx >= n) { *diag = 1.0; return; }
}

__global__ void k_mask_diagonal_poses(DeviceProblem P, const uint8_t* mask_used) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P.n_poses) return;
  const int o = P.pose_opt[p];
  if (o < 0 || !mask_used[p]) return;
  const uint32_t fr = P.pose_free[p];
  for (int k = 0; k < P.D; ++k) {
    if (!((fr >> k) & 1u)) {
      const int idx = o * P.D + k;
      const int tI = idx / P.T;
      P.S_tiles[(size_t)P.tile_map[tI * P.nT + tI] * P.T * P.T + (size_t)(idx % P.T) * P.T + (idx % P.T)] = 1e6;
    }
  }
}

// ---------------------------------------------------------------------------------------
// small vector kernels for the dogleg step control
__global__ void k_dot(const double* a, const double* b, size_t n, double* out) {
  double acc = 0.0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc += a[i] * b[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red_add(out, acc);
}
// out = alpha * x + beta * y  (y may be null)
__global__ void k_axpby(double* out, double alpha, const double* x, double beta, const double* y, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = alpha * x[i] + (y ? beta * y[i] : 0.0);
}
// gather b_l (indexed by landmark id) into opt-id order (rhs_l_)
template <int L>
__global__ void k_compact_bl(DeviceProblem P, double* out) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  const int o = P.lm_opt[lm];
  if (o < 0) return;
#pragma unroll
  for (int k = 0; k < L; ++k) out[(size_t)o * L + k] = P.lm_bl[(size_t)lm * L + k];
}

// ---------------------------------------------------------------------------------------
// launchers
static int g_sms = 148;
void set_sm_count(int n) { g_sms = n > 0 ? n : 148; }

static inline int grid_for(size_t work_items, int block, int max_per_sm) {
  size_t b = (work_items + block - 1) / block;
  size_t cap = (size_t)g_sms * max_per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

#define DISPATCH_LG(L, G, CALL)                                                     \
  do {                                                                              \
    if (L == 1) {                                                                   \
      if (G == 4) { CALL(1, 4); } else if (G == 8) { CALL(1, 8); } else if (G == 16) { CALL(1, 16); } else { CALL(1, 32); } \
    } else {                                                                        \
      if (G == 4) { CALL(3, 4); } else if (G == 8) { CALL(3, 8); } else if (G == 16) { CALL(3, 16); } else { CALL(3, 32); } \
    }                                                                               \
  } while (0)

void launch_pose_frames(const DeviceProblem& P, cudaStream_t st) {
  const int n = P.n_poses * P.n_cams;
  if (n == 0) return;
  k_pose_frames<<<(n + 127) / 128, 128, 0, st>>>(P);
}
void launch_prepare_xs(const DeviceProblem& P, const double* lm_xw, cudaStream_t st) {
  if (P.n_lms == 0) return;
  k_prepare_xs<<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, lm_xw);
}
void launch_finish_xw(const DeviceProblem& P, double* lm_xw, cudaStream_t st) {
  if (P.n_lms == 0) return;
  k_finish_xw<<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, lm_xw);
}
void launch_residual(const DeviceProblem& P, int G, cudaStream_t st) {
  if (P.n_obs == 0) return;
  const int grid = grid_for((size_t)P.n_lms * G, 256, 8);
#define CALL(LL, GG) k_residual<LL, GG><<<grid, 256, 0, st>>>(P)
  DISPATCH_LG(P.L, G, CALL);
#undef CALL
}
void launch_accumulate(const DeviceProblem& P, int G, cudaStream_t st) {
  if (P.n_obs == 0) return;
  const int grid = grid_for((size_t)P.n_lms * G, 128, 12);
#define CALL(LL, GG) k_accumulate<LL, GG><<<grid, 128, 0, st>>>(P)
  DISPATCH_LG(P.L, G, CALL);
#undef CALL
}
void launch_schur(const DeviceProblem& P, cudaStream_t st) {
  if (P.n_obs == 0 || P.n_active_lms == 0 || P.n_active_poses == 0) return;
  const int grid = grid_for((size_t)P.n_lms * 32, 128, 8);
  if (P.L == 1) k_schur<1><<<grid, 128, 0, st>>>(P);
  else k_schur<3><<<grid, 128, 0, st>>>(P);
}
void launch_sd_denominator(const DeviceProblem& P, int G, cudaStream_t st) {
  if (P.n_obs == 0) return;
  const int grid = grid_for((size_t)P.n_lms * G, 128, 12);
#define CALL(LL, GG) k_sd_denominator<LL, GG><<<grid, 128, 0, st>>>(P)
  DISPATCH_LG(P.L, G, CALL);
#undef CALL
}
void launch_back_substitute(const DeviceProblem& P, int G, const double* delta_p, double* delta_l, cudaStream_t st) {
  if (P.n_active_lms == 0) return;
  const int grid = grid_for((size_t)P.n_lms * G, 256, 8);
#define CALL(LL, GG) k_back_substitute<LL, GG><<<grid, 256, 0, st>>>(P, delta_p, delta_l)
  DISPATCH_LG(P.L, G, CALL);
#undef CALL
}
void launch_evaluate(const DeviceProblem& P, int G, double* out_sum, double* res_out, double* maha_out, cudaStream_t st) {
  if (P.n_obs == 0) return;
  cudaMemsetAsync(P.lm_outliers, 0, sizeof(uint32_t) * P.n_lms, st);
  const int grid = grid_for((size_t)P.n_lms * G, 256, 8);
#define CALL(LL, GG) k_evaluate<LL, GG><<<grid, 256, 0, st>>>(P, out_sum, res_out, maha_out)
  DISPATCH_LG(P.L, G, CALL);
#undef CALL
}
void launch_update_poses(const DeviceProblem& P, const double* delta_p, double coef, cudaStream_t st) {
  if (P.n_poses == 0) return;
  k_update_poses<<<(P.n_poses + 127) / 128, 128, 0, st>>>(P, delta_p, coef);
}
void launch_update_landmarks(const DeviceProblem& P, const double* delta_l, double coef, cudaStream_t st) {
  if (P.n_lms == 0 || P.L == 0) return;
  if (P.L == 1) k_update_landmarks<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, delta_l, coef);
  else k_update_landmarks<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, delta_l, coef);
}
void launch_mask_diagonal(const DeviceProblem& P, const uint8_t* mask_used, int global_flag, cudaStream_t st) {
  const int n_pad = P.nT * P.T;
  if (n_pad == 0) return;
  k_mask_diagonal<<<(n_pad + 255) / 256, 256, 0, st>>>(P, mask_used, global_flag);
  if (global_flag) k_mask_diagonal_poses<<<(P.n_poses + 127) / 128, 128, 0, st>>>(P, mask_used);
}
void launch_dot(const double* a, const double* b, size_t n, double* out, cudaStream_t st) {
  if (n == 0) return;
  k_dot<<<grid_for(n, 256, 4), 256, 0, st>>>(a, b, n, out);
}
void launch_axpby(double* out, double alpha, const double* x, double beta, const double* y, size_t n, cudaStream_t st) {
  if (n == 0) return;
  k_axpby<<<grid_for(n, 256, 4), 256, 0, st>>>(out, alpha, x, beta, y, n);
}
void launch_compact_bl(const DeviceProblem& P, double* out, cudaStream_t st) {
  if (P.n_lms == 0 || P.L == 0) return;
  if (P.L == 1) k_compact_bl<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, out);
  else k_compact_bl<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, out);
}

}  // namespace bagpu
