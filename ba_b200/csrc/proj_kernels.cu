// proj_kernels.cu — kernels and launchers for the projection part of the hot path.
// See proj_kernels.cuh for the reference lines each kernel replaces.
#include <mutex>
#include <unordered_map>
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
// One thread per observation of the landmark-major stream (fully coalesced, every lane busy: with one lane group per
// landmark only 13 of 32 lanes were active on a BAL-like track-length distribution).
template <int L>
__global__ void __launch_bounds__(256) k_residual(DeviceProblem P) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n_obs; i += gridDim.x * blockDim.x) {
    LmCtx<L> c;
    load_lm<L>(P, (int)P.obs_lm[i], c);
    double r[2];
    obs_residual<L>(P, c, i, r);
    P.obs_maha0[i] = (r[0] * r[0] + r[1] * r[1]) * P.obs_w0[i];
  }
}

// pass 2 (Jacobians, weights, landmark blocks, pose-major records) and the pose-pair-major
// reduction into S live in schur_kernels.cu.

// ---------------------------------------------------------------------------------------
// steepest-descent denominator |J_p b_p + J_l b_l|^2 over projection residuals (cpp:879-906)
// Every projection Jacobian is B times something (ba_gpu.h), so J_p b_p + J_l b_l = B v with
//   v = K_meas b_meas - K_ref b_ref - (b_l  or  u b_l)     and     w |B v|^2 = v^T G v
// from the G record of the last build: no camera model here.
// per-pose operands of K_pose delta (proj_kernels.cuh, k_times_pack) for the vector `delta` over the optimised poses
__global__ void k_pose_delta_pack(DeviceProblem P, const double* __restrict__ delta, double* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P.n_poses) return;
  const int o = P.pose_opt[p];
  double x[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, xw[3] = {0.0, 0.0, 0.0};
  if (o >= 0) {
    const uint32_t fr = P.pose_free[p];
#pragma unroll
    for (int a = 0; a < 6; ++a) x[a] = ((fr >> a) & 1u) ? delta[(size_t)o * P.D + a] : 0.0;
    mat3_vec(P.Rwp + 9 * (size_t)p, x + 3, xw);
  } else {
    x[0] = __longlong_as_double(0x7ff8000000000000LL);
  }
  const double* tp = P.pose_t + 3 * (size_t)p;
  const double* tw = P.Tws + (size_t)p * P.n_cams * 12 + 9;
  double* q = out + (size_t)p * kPosePack;
  q[0] = x[0]; q[1] = x[1]; q[2] = x[2]; q[3] = xw[0]; q[4] = xw[1]; q[5] = xw[2];
  q[6] = tp[0]; q[7] = tp[1]; q[8] = tp[2]; q[9] = tw[0]; q[10] = tw[1]; q[11] = tw[2];
}

template <int L>
__global__ void __launch_bounds__(128) k_sd_denominator(DeviceProblem P) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n_obs; i += gridDim.x * blockDim.x) {
    const int lm = (int)P.obs_lm[i];
    LmRecCtx c;
    load_lm_rec<L>(P, lm, c);
    const int m = (int)P.obs_pose[i];
    const bool diff = (m != c.ref_pose);
    double v[3], tw[3];
    k_times_pack(P.kd_pack + (size_t)m * kPosePack, c.X, c.rho, v, tw);
    if (!diff) { v[0] = 0.0; v[1] = 0.0; v[2] = 0.0; }
    if (L == 1 && diff) {
      double y[3], tr[3];
      if (k_times_pack(P.kd_pack + (size_t)c.ref_pose * kPosePack, c.X, c.rho, y, tr)) { v[0] -= y[0]; v[1] -= y[1]; v[2] -= y[2]; }
    }
    if (P.lm_opt[lm] >= 0) {
      if (L == 3) {
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] -= P.lm_bl[(size_t)lm * 3 + k];
      } else {
        if (P.n_cams > 1 && P.obs_cam) {
          const double* t2 = P.Tws + ((size_t)m * P.n_cams + (int)P.obs_cam[i]) * 12 + 9;
          tw[0] = t2[0]; tw[1] = t2[1]; tw[2] = t2[2];
        }
        const double bl = P.lm_bl[lm];
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] -= (c.tws_r[k] - tw[k]) * bl;
      }
    }
    double G[6], gv[3];
    load_G(P, i, G);
    sym3_vec(G, v, gv);
    acc[0] += v[0] * gv[0] + v[1] * gv[1] + v[2] * gv[2];
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->sd_den_proj, false);
}

// ---------------------------------------------------------------------------------------
// The rows of j_pr_ / r_pr_ / j_l_ the reference inserts per residual (cpp:1636-1642, 1795-1796), written at the
// residual's id for the dump (cpp:607-620).  The build never forms them: this kernel exists for the interchange files
// only.  out (28 + 2 L doubles per residual): r sqrt(w) (2) | dz_dx_meas sqrt(w) (12) | dz_dx_ref sqrt(w) (12) |
// dz_dlm sqrt(w) (2 L) | meas opt id, ref opt id as doubles (-1: no block)
template <int L>
__global__ void __launch_bounds__(128) k_jacobian_rows(DeviceProblem P, double* __restrict__ out) {
  constexpr int W = 28 + 2 * L;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n_obs; i += gridDim.x * blockDim.x) {
    const int lm = (int)P.obs_lm[i];
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    ObsJac<L> o;
    obs_jacobians<L>(P, c, i, o);
    const double sw = sqrt(P.obs_w[i]);
    double* q = out + (size_t)P.obs_perm[i] * W;
    q[0] = o.r[0] * sw; q[1] = o.r[1] * sw;
#pragma unroll
    for (int k = 0; k < 12; ++k) q[2 + k] = (o.pm >= 0) ? o.Jm[k] * sw : 0.0;
    double Jr[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) Jr[k] = 0.0;
    if (L == 1 && o.ref_on) ref_jacobian<L>(c, o.Bp, Jr);
#pragma unroll
    for (int k = 0; k < 12; ++k) q[14 + k] = Jr[k] * sw;
#pragma unroll
    for (int k = 0; k < 2 * L; ++k) q[26 + k] = (c.lm_opt >= 0) ? o.Jl[k] * sw : 0.0;
    q[26 + 2 * L] = (double)o.pm;
    q[27 + 2 * L] = (L == 1 && o.ref_on) ? (double)c.ref_opt : -1.0;
  }
}

// ---------------------------------------------------------------------------------------
// back-substitution: delta_l = V^-1 (b_l - W^T delta_p) = dhat - V^-1 sum_a W_a^T delta_p(a)   (cpp:709-744).
// W_a^T delta = w J_l^T (J_m delta) is re-evaluated from the observation (nothing per observation is stored);
// V^-1 and dhat come from the landmark's record.
// share of one observation: w J_l^T J_m delta_p(a)
template <int L>
BA_D void back_term(const DeviceProblem& P, const LmRecCtx& c, int i, double* acc) {
  // W_a^T delta = w J_l^T J_m delta = -G (K_m delta)   (LmSize 1: -u^T G K_m delta), from the G record of the last build and
  // the pose's pack for delta_p (launch_back_substitute)
  const int m = (int)P.obs_pose[i];
  if (m == c.ref_pose) return;
  double y[3], tw[3];
  if (!k_times_pack(P.kd_pack + (size_t)m * kPosePack, c.X, c.rho, y, tw)) return;
  if (L == 3) {
    double G[6], gy[3];
    load_G(P, i, G);
    sym3_vec(G, y, gy);
#pragma unroll
    for (int k = 0; k < 3; ++k) acc[k] -= gy[k];
  } else {
    // u^T G y = (G u) . y: the 32-byte gv record of the observation instead of G and the two sensor positions
    double g0, g1, g2, vi;
    ld4(P.obs_gv + (size_t)i * kObsGv, g0, g1, g2, vi);
    (void)vi;
    acc[0] -= g0 * y[0] + g1 * y[1] + g2 * y[2];
  }
}
// landmark part and write-out, by one lane holding the landmark's sum
template <int L>
BA_D void back_finish(const DeviceProblem& P, int lm, int lo, double* acc, double* delta_l) {
  const double* r = P.lm_rec + (size_t)lm * kLmRec;
  if (L == 1) {
    const int rp = (int)P.lm_ref_pose[lm];
    double k[3], tw[3];     // W_ref^T delta = gbar^T K_ref delta
    if (k_times_pack(P.kd_pack + (size_t)rp * kPosePack, r, r[3], k, tw)) acc[0] += r[9] * k[0] + r[10] * k[1] + r[11] * k[2];
    delta_l[lo] = r[8] - r[7] * acc[0];
  } else {
    const double v00 = r[4], v01 = r[5], v02 = r[6], v11 = r[7], v12 = r[8], v22 = r[9];
    delta_l[(size_t)lo * L + 0] = r[10] - (v00 * acc[0] + v01 * acc[1] + v02 * acc[2]);
    delta_l[(size_t)lo * L + 1] = r[11] - (v01 * acc[0] + v11 * acc[1] + v12 * acc[2]);
    delta_l[(size_t)lo * L + 2] = r[12] - (v02 * acc[0] + v12 * acc[1] + v22 * acc[2]);
  }
}

// packed tasks (see k_linearize_packed): one observation per lane, segmented sum over the lanes of a landmark
template <int L>
__global__ void __launch_bounds__(256) k_back_substitute_packed(DeviceProblem P, const uint2* __restrict__ tasks, int n_tasks, double* delta_l) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (int t = warp; t < n_tasks; t += n_warps) {
    const uint2 tk = tasks[t];
    const int s = (int)tk.x, n = (int)tk.y;
    const bool on = lane < n;
    const int i = s + (on ? lane : 0);
    const int lm = (int)P.obs_lm[i];
    const int lm_prev = __shfl_up_sync(0xffffffffu, lm, 1);
    const bool head = on && (lane == 0 || lm != lm_prev);
    const unsigned heads = __ballot_sync(0xffffffffu, head);
    const unsigned above = (lane == 31) ? 0u : (heads & (0xffffffffu << (lane + 1)));
    const int seg_end = above ? (__ffs(above) - 1) : n;
    const int lo = P.lm_opt[lm];
    double acc[L];
#pragma unroll
    for (int k = 0; k < L; ++k) acc[k] = 0.0;
    if (on && lo >= 0) {
      LmRecCtx c;
      load_lm_rec<L, false>(P, lm, c);
      back_term<L>(P, c, i, acc);
    }
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const bool take = on && (lane + d < seg_end);
#pragma unroll
      for (int k = 0; k < L; ++k) {
        const double other = __shfl_down_sync(0xffffffffu, acc[k], d);
        if (take) acc[k] += other;
      }
    }
    if (head && lo >= 0) back_finish<L>(P, lm, lo, acc, delta_l);
  }
}

// landmarks with more than 32 observations: one warp per landmark
template <int L>
__global__ void __launch_bounds__(256) k_back_substitute_long(DeviceProblem P, const uint32_t* __restrict__ lm_list, int n_list, double* delta_l) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (int q = warp; q < n_list; q += n_warps) {
    const int lm = (int)lm_list[q];
    const int lo = P.lm_opt[lm];
    if (lo < 0) continue;
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    double acc[L];
#pragma unroll
    for (int k = 0; k < L; ++k) acc[k] = 0.0;
    LmRecCtx c;
    load_lm_rec<L, false>(P, lm, c);
    for (int i = s + lane; i < e; i += 32) back_term<L>(P, c, i, acc);
#pragma unroll
    for (int k = 0; k < L; ++k) acc[k] = warp_sum(acc[k]);
    if (lane == 0) back_finish<L>(P, lm, lo, acc, delta_l);
  }
}

// ---------------------------------------------------------------------------------------
// residual-only evaluation with the current Huber weights (cpp:148-188)
template <int L>
__global__ void __launch_bounds__(256) k_evaluate(DeviceProblem P, double* out_sum, double* res_out, double* maha_out, int write_maha0, int count_outliers) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n_obs; i += gridDim.x * blockDim.x) {
    const int lm = (int)P.obs_lm[i];
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    double r[2];
    obs_residual<L>(P, c, i, r);
    const double n2 = r[0] * r[0] + r[1] * r[1];
    const double m = n2 * P.obs_w[i];
    acc[0] += m;
    // the next BuildProblem starts from this state if the step is accepted: its median input, for free
    if (write_maha0) P.obs_maha0[i] = n2 * P.obs_w0[i];
    if (count_outliers && sqrt(n2) > P.outlier_threshold) atomicAdd(&P.lm_outliers[lm], 1u);   // cpp:184-186
    if (res_out) {
      const uint32_t id = P.obs_perm[i];
      res_out[2 * (size_t)id] = r[0]; res_out[2 * (size_t)id + 1] = r[1];
      maha_out[id] = m;
    }
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, out_sum, true);
}

// ---------------------------------------------------------------------------------------
// ApplyUpdate (cpp:84-139): poses and landmarks, coefficient `coef`
// bk.q != nullptr: the state before the update goes to the backup arrays on the way (poses_copy / landmarks_copy,
// cpp:1020-1023: one kernel instead of seven device-to-device copies)
__global__ void k_update_poses(DeviceProblem P, const double* delta_p, double coef, StateBackup bk) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P.n_poses) return;
  if (bk.q) {
    for (int k = 0; k < 4; ++k) bk.q[4 * (size_t)p + k] = P.pose_q[4 * (size_t)p + k];
    for (int k = 0; k < 3; ++k) { bk.t[3 * (size_t)p + k] = P.pose_t[3 * (size_t)p + k]; bk.v[3 * (size_t)p + k] = P.pose_v[3 * (size_t)p + k]; }
    for (int k = 0; k < 6; ++k) bk.b[6 * (size_t)p + k] = P.pose_b[6 * (size_t)p + k];
  }
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
__global__ void k_update_landmarks(DeviceProblem P, const double* delta_l, double coef, StateBackup bk) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  double* x = P.lm_x + 4 * (size_t)lm;
  if (bk.lm_x) {
    reinterpret_cast<double4*>(bk.lm_x)[lm] = *reinterpret_cast<const double4*>(x);
    bk.rel[lm] = P.lm_reliable[lm];
    bk.outliers[lm] = P.lm_outliers[lm];
    P.lm_outliers[lm] = 0;      // the evaluation that follows recounts them (launch_evaluate, count_outliers = 2)
  }
  const int o = P.lm_opt[lm];
  if (o < 0) return;
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
  if (idx >= n) *s_diag(P, idx) = 1.0;
}

__global__ void k_mask_diagonal_poses(DeviceProblem P, const uint8_t* mask_used) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P.n_poses) return;
  const int o = P.pose_opt[p];
  if (o < 0 || !mask_used[p]) return;
  const uint32_t fr = P.pose_free[p];
  for (int k = 0; k < P.D; ++k) {
    if (!((fr >> k) & 1u)) {
      *s_diag(P, o * P.D + k) = 1e6;
    }
  }
}

// ---------------------------------------------------------------------------------------
// small vector kernels for the dogleg step control
__global__ void k_dot(const double* a, const double* b, size_t n, double* out, double* partial, unsigned int* counter) {
  double acc[1] = {0.0};
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc[0] += a[i] * b[i];
  grid_sum_fixed<1>(acc, partial, counter, out, false);
}
// out = alpha * x + beta * y  (y may be null)
__global__ void k_axpby(double* out, double alpha, const double* x, double beta, const double* y, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = alpha * x[i] + (y ? beta * y[i] : 0.0);
}
// Dogleg blend (cpp:991-1017), poses and landmarks in one pass each (index space [0, np) + [np, np + nl)):
//   first half   tmp = gn - f sd, out = { |tmp_p|^2, tmp_p . sd_p, |tmp_l|^2, tmp_l . sd_l }
//   second half  dl = beta tmp + f sd, out = { |dl_p|^2, |dl_l|^2 }
// (two launches and one host round trip instead of ten launches and three)
__global__ void __launch_bounds__(256) k_blend_diff(double* __restrict__ tmp_p, const double* __restrict__ gn_p, const double* __restrict__ sd_p, size_t np,
                                                    double* __restrict__ tmp_l, const double* __restrict__ gn_l, const double* __restrict__ sd_l, size_t nl,
                                                    double f, double* out, double* partial, unsigned int* counter) {
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < np + nl; i += (size_t)gridDim.x * blockDim.x) {
    if (i < np) { const double s = sd_p[i], t = gn_p[i] - f * s; tmp_p[i] = t; acc[0] += t * t; acc[1] += t * s; }
    else { const size_t j = i - np; const double s = sd_l[j], t = gn_l[j] - f * s; tmp_l[j] = t; acc[2] += t * t; acc[3] += t * s; }
  }
  grid_sum_fixed<4>(acc, partial, counter, out, false);
}
__global__ void __launch_bounds__(256) k_blend_step(double* __restrict__ dl_p, const double* __restrict__ tmp_p, const double* __restrict__ sd_p, size_t np,
                                                    double* __restrict__ dl_l, const double* __restrict__ tmp_l, const double* __restrict__ sd_l, size_t nl,
                                                    double beta, double f, double* out, double* partial, unsigned int* counter) {
  double acc[2] = {0.0, 0.0};
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < np + nl; i += (size_t)gridDim.x * blockDim.x) {
    if (i < np) { const double d = beta * tmp_p[i] + f * sd_p[i]; dl_p[i] = d; acc[0] += d * d; }
    else { const size_t j = i - np; const double d = beta * tmp_l[j] + f * sd_l[j]; dl_l[j] = d; acc[1] += d * d; }
  }
  grid_sum_fixed<2>(acc, partial, counter, out, false);
}
// tile-wise permutation of a padded pose vector: forward (natural -> elimination order) or back
__global__ void k_permute_tiles(double* dst, const double* src, const int* perm, int T, int nT, int forward) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nT * T) return;
  const int t = idx / T, o = idx - t * T;
  if (forward) dst[(size_t)perm[t] * T + o] = src[idx];
  else dst[idx] = src[(size_t)perm[t] * T + o];
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

int resident_ctas_per_sm(const void* kernel, int block, size_t smem) {
  static std::mutex mu;
  static std::unordered_map<const void*, int> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(kernel);
  if (it != cache.end()) return it->second;
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, block, smem) != cudaSuccess || n < 1) n = 1;
  cache[kernel] = n;
  return n;
}

// one wave of resident CTAs (at most), never more than the fixed-order reductions have slots for
template <typename K>
static inline int grid_for(K kernel, size_t work_items, int block) {
  size_t b = (work_items + block - 1) / block;
  size_t cap = (size_t)g_sms * resident_ctas_per_sm(reinterpret_cast<const void*>(kernel), block);
  if (cap > (size_t)kRedMaxGrid) cap = kRedMaxGrid;
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
  (void)G;
  if (P.L == 1) k_residual<1><<<grid_for(k_residual<1>, (size_t)P.n_obs, 256), 256, 0, st>>>(P);
  else k_residual<3><<<grid_for(k_residual<3>, (size_t)P.n_obs, 256), 256, 0, st>>>(P);
}
void launch_jacobian_rows(const DeviceProblem& P, double* out, cudaStream_t st) {
  if (P.n_obs == 0 || P.L == 0) return;
  if (P.L == 1) k_jacobian_rows<1><<<grid_for(k_jacobian_rows<1>, (size_t)P.n_obs, 128), 128, 0, st>>>(P, out);
  else k_jacobian_rows<3><<<grid_for(k_jacobian_rows<3>, (size_t)P.n_obs, 128), 128, 0, st>>>(P, out);
}
void launch_sd_denominator(const DeviceProblem& P, int G, cudaStream_t st) {
  if (P.n_obs == 0) return;
  (void)G;
  k_pose_delta_pack<<<(P.n_poses + 127) / 128, 128, 0, st>>>(P, P.b_p, P.kd_pack);
  if (P.L == 1) k_sd_denominator<1><<<grid_for(k_sd_denominator<1>, (size_t)P.n_obs, 128), 128, 0, st>>>(P);
  else k_sd_denominator<3><<<grid_for(k_sd_denominator<3>, (size_t)P.n_obs, 128), 128, 0, st>>>(P);
}
void launch_back_substitute(const DeviceProblem& P, const uint2* tasks, int n_tasks, const uint32_t* long_lms, int n_long, const double* delta_p,
                            double* delta_l, cudaStream_t st) {
  if (P.n_active_lms == 0) return;
  k_pose_delta_pack<<<(P.n_poses + 127) / 128, 128, 0, st>>>(P, delta_p, P.kd_pack);
  if (n_tasks > 0) {
    if (P.L == 1) k_back_substitute_packed<1><<<grid_for(k_back_substitute_packed<1>, (size_t)n_tasks * 32, 256), 256, 0, st>>>(P, tasks, n_tasks, delta_l);
    else k_back_substitute_packed<3><<<grid_for(k_back_substitute_packed<3>, (size_t)n_tasks * 32, 256), 256, 0, st>>>(P, tasks, n_tasks, delta_l);
  }
  if (n_long > 0) {
    if (P.L == 1) k_back_substitute_long<1><<<grid_for(k_back_substitute_long<1>, (size_t)n_long * 32, 256), 256, 0, st>>>(P, long_lms, n_long, delta_l);
    else k_back_substitute_long<3><<<grid_for(k_back_substitute_long<3>, (size_t)n_long * 32, 256), 256, 0, st>>>(P, long_lms, n_long, delta_l);
  }
}
// count_outliers: EvaluateResiduals resets and recounts num_outlier_residuals (cpp:148-186); the read-out passes after
// Solve() (per-residual export, conditioning summary) must leave the counts alone (on several GPUs they have been
// merged over the ranks by then)
void launch_evaluate(const DeviceProblem& P, int G, double* out_sum, double* res_out, double* maha_out, int write_maha0, int count_outliers,
                     cudaStream_t st) {
  if (P.n_obs == 0) return;
  if (count_outliers == 1) cudaMemsetAsync(P.lm_outliers, 0, sizeof(uint32_t) * P.n_lms, st);   // 2: cleared by k_update_landmarks
  (void)G;
  if (P.L == 1) k_evaluate<1><<<grid_for(k_evaluate<1>, (size_t)P.n_obs, 256), 256, 0, st>>>(P, out_sum, res_out, maha_out, write_maha0, count_outliers);
  else k_evaluate<3><<<grid_for(k_evaluate<3>, (size_t)P.n_obs, 256), 256, 0, st>>>(P, out_sum, res_out, maha_out, write_maha0, count_outliers);
}
void launch_update_poses(const DeviceProblem& P, const double* delta_p, double coef, cudaStream_t st, const StateBackup* bk) {
  if (P.n_poses == 0) return;
  StateBackup b = {};
  if (bk) b = *bk;
  k_update_poses<<<(P.n_poses + 127) / 128, 128, 0, st>>>(P, delta_p, coef, b);
}
void launch_update_landmarks(const DeviceProblem& P, const double* delta_l, double coef, cudaStream_t st, const StateBackup* bk) {
  if (P.n_lms == 0 || P.L == 0) return;
  StateBackup b = {};
  if (bk) b = *bk;
  if (P.L == 1) k_update_landmarks<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, delta_l, coef, b);
  else k_update_landmarks<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, delta_l, coef, b);
}
void launch_mask_diagonal(const DeviceProblem& P, const uint8_t* mask_used, int global_flag, cudaStream_t st) {
  const int n_pad = P.nT * P.T;
  if (n_pad == 0) return;
  k_mask_diagonal<<<(n_pad + 255) / 256, 256, 0, st>>>(P, mask_used, global_flag);
  if (global_flag) k_mask_diagonal_poses<<<(P.n_poses + 127) / 128, 128, 0, st>>>(P, mask_used);
}
void launch_dot(const DeviceProblem& P, const double* a, const double* b, size_t n, double* out, cudaStream_t st) {
  if (n == 0) return;
  k_dot<<<grid_for(k_dot, n, 256), 256, 0, st>>>(a, b, n, out, P.red_partial, P.red_counter);
}
void launch_axpby(double* out, double alpha, const double* x, double beta, const double* y, size_t n, cudaStream_t st) {
  if (n == 0) return;
  k_axpby<<<grid_for(k_axpby, n, 256), 256, 0, st>>>(out, alpha, x, beta, y, n);
}
void launch_blend_diff(const DeviceProblem& P, double* tmp_p, const double* gn_p, const double* sd_p, size_t np, double* tmp_l, const double* gn_l,
                       const double* sd_l, size_t nl, double f, double* out4, cudaStream_t st) {
  k_blend_diff<<<grid_for(k_blend_diff, np + nl, 256), 256, 0, st>>>(tmp_p, gn_p, sd_p, np, tmp_l, gn_l, sd_l, nl, f, out4, P.red_partial, P.red_counter);
}
void launch_blend_step(const DeviceProblem& P, double* dl_p, const double* tmp_p, const double* sd_p, size_t np, double* dl_l, const double* tmp_l,
                       const double* sd_l, size_t nl, double beta, double f, double* out2, cudaStream_t st) {
  k_blend_step<<<grid_for(k_blend_step, np + nl, 256), 256, 0, st>>>(dl_p, tmp_p, sd_p, np, dl_l, tmp_l, sd_l, nl, beta, f, out2, P.red_partial, P.red_counter);
}
void launch_permute_tiles(double* dst, const double* src, const int* perm, int T, int nT, int forward, cudaStream_t st) {
  if (nT * T == 0) return;
  k_permute_tiles<<<(nT * T + 255) / 256, 256, 0, st>>>(dst, src, perm, T, nT, forward);
}
void launch_compact_bl(const DeviceProblem& P, double* out, cudaStream_t st) {
  if (P.n_lms == 0 || P.L == 0) return;
  if (P.L == 1) k_compact_bl<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, out);
  else k_compact_bl<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P, out);
}

}  // namespace bagpu
