// proj_kernels.cuh — projection-residual kernels of the hot path (sm_100a, FP64).
//
// Replaces, per outer iteration of ba::BundleAdjuster::Solve:
//   * ParallelProjectionResiduals::operator()      reference include/ba/parallel_algos.h:35-152
//   * Huber re-weighting                           reference src/BundleAdjuster.cpp:1353-1388
//   * J^T J / J^T r accumulation into U, b_p       reference src/BundleAdjuster.cpp:337-354
//   * per-landmark V, b_l, V^-1                    reference src/BundleAdjuster.cpp:409-443
//   * W = J_pr^T J_l                               reference src/BundleAdjuster.cpp:448-455
//   * Schur complement S = U - W V^-1 W^T, rhs     reference src/BundleAdjuster.cpp:459-484
//   * steepest-descent denominator                 reference src/BundleAdjuster.cpp:879-910
//   * back-substitution                            reference src/BundleAdjuster.cpp:709-744
//   * residual-only evaluation                     reference src/BundleAdjuster.cpp:148-188
// The Jacobians are the algebraically collapsed closed forms of the reference's
// 2x4 * 4x7 * 7x7 * 7x6 chain; J is never stored, every kernel that needs a Jacobian
// re-evaluates it from the 40-byte observation, the landmark and the pose frame.
//
// Work decomposition: observations are sorted by landmark.  The landmark-major kernels take one
// observation per lane (warp tasks = runs of whole landmarks with at most 32 observations, packed at
// set-up); per-landmark sums are segmented warp reductions.  The pose-pair-major kernel
// (schur_kernels.cu) owns each S block exclusively: no atomics anywhere in the build.
#pragma once
#include "ba_gpu.h"
#include "device_math.cuh"

namespace bagpu {

template <int L>
struct LmCtx {
  double X[3];      // world point (LmSize 3: x_w; LmSize 1: R_ws_r ray + rho t_ws_r)
  double rho;       // homogeneous scale (x_w[3] or inverse depth)
  double tws_r[3];  // LmSize 1: reference sensor position in the world
  double Ar[9];     // LmSize 1: R_wp_r [Y]x, Y = R_vs_r ray + rho t_vs_r
  int ref_pose, ref_opt, lm_opt;
  uint32_t ref_free;
};

template <int L>
BA_D void load_lm(const DeviceProblem& P, int lm, LmCtx<L>& c) {
  const double4 x = *reinterpret_cast<const double4*>(P.lm_x + 4 * (size_t)lm);
  c.ref_pose = (int)P.lm_ref_pose[lm];
  c.lm_opt = P.lm_opt[lm];
  c.rho = x.w;
  if (L == 1) {
    const int rc = (int)P.lm_ref_cam[lm];
    const double* Tws = P.Tws + ((size_t)c.ref_pose * P.n_cams + rc) * 12;
    const double ray[3] = {x.x, x.y, x.z};
    double Xw[3];
    mat3_vec(Tws, ray, Xw);
    c.tws_r[0] = Tws[9]; c.tws_r[1] = Tws[10]; c.tws_r[2] = Tws[11];
    c.X[0] = Xw[0] + x.w * c.tws_r[0]; c.X[1] = Xw[1] + x.w * c.tws_r[1]; c.X[2] = Xw[2] + x.w * c.tws_r[2];
    double Y[3];
    mat3_vec(P.cam_Rvs + 9 * rc, ray, Y);
    Y[0] += x.w * P.cam_tvs[3 * rc]; Y[1] += x.w * P.cam_tvs[3 * rc + 1]; Y[2] += x.w * P.cam_tvs[3 * rc + 2];
    mat3_skew(P.Rwp + 9 * (size_t)c.ref_pose, Y, c.Ar);
    c.ref_opt = P.pose_opt[c.ref_pose];
    c.ref_free = c.ref_opt >= 0 ? P.pose_free[c.ref_pose] : 0u;
  } else {
    c.X[0] = x.x; c.X[1] = x.y; c.X[2] = x.z;
    c.ref_opt = -1; c.ref_free = 0;
  }
}

// a pose frame [R (9, row-major) | t (3)] in six 16-byte loads (frames are 96 bytes apart)
BA_D void load_frame(const double* __restrict__ T, double* f) {
#pragma unroll
  for (int k = 0; k < 6; ++k) { const double2 v = reinterpret_cast<const double2*>(T)[k]; f[2 * k] = v.x; f[2 * k + 1] = v.y; }
}

template <int L>
struct ObsJac {
  double r[2];
  double Jl[2 * L];   // row-major 2 x L
  double Jm[12];      // row-major 2 x 6 (measurement pose)
  double Bp[6];       // row-major 2 x 3: d project / d X_world = Jp R_sw (the reference-pose Jacobian is Bp [-rho I | Ar])
  int pm;             // opt id of the measurement pose if it takes part (active, diff_poses), else -1
  bool ref_on;        // reference pose takes part (LmSize 1, active, diff_poses)
};

// residual only (parallel_algos.h:59-64, src/BundleAdjuster.cpp:159-173)
template <int L>
BA_D void obs_residual(const DeviceProblem& P, const LmCtx<L>& c, int obs, double* r) {
  const int m = (int)P.obs_pose[obs];
  const int cam = P.obs_cam ? (int)P.obs_cam[obs] : 0;
  double Tsw[12];
  load_frame(P.Tsw + ((size_t)m * P.n_cams + cam) * 12, Tsw);
  double Xs[3];
  mat3_vec(Tsw, c.X, Xs);
  Xs[0] += c.rho * Tsw[9]; Xs[1] += c.rho * Tsw[10]; Xs[2] += c.rho * Tsw[11];
  const double* K = P.cam_params + 4 * cam;
  const double iz = 1.0 / Xs[2];
  const double2 z = P.obs_z[obs];
  r[0] = z.x - (K[0] * Xs[0] * iz + K[2]);
  r[1] = z.y - (K[1] * Xs[1] * iz + K[3]);
}

// residual + closed-form Jacobians
template <int L>
BA_D void obs_jacobians(const DeviceProblem& P, const LmCtx<L>& c, int obs, ObsJac<L>& o, int m_known = -1) {
  const int m = m_known >= 0 ? m_known : (int)P.obs_pose[obs];
  const int cam = P.obs_cam ? (int)P.obs_cam[obs] : 0;
  double Tsw[12];
  load_frame(P.Tsw + ((size_t)m * P.n_cams + cam) * 12, Tsw);
  double R[9], t[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = Tsw[i];
  t[0] = Tsw[9]; t[1] = Tsw[10]; t[2] = Tsw[11];
  double Xs[3];
  mat3_vec(R, c.X, Xs);
  Xs[0] += c.rho * t[0]; Xs[1] += c.rho * t[1]; Xs[2] += c.rho * t[2];
  const double* K = P.cam_params + 4 * cam;
  const double iz = 1.0 / Xs[2];
  const double2 z = P.obs_z[obs];
  o.r[0] = z.x - (K[0] * Xs[0] * iz + K[2]);
  o.r[1] = z.y - (K[1] * Xs[1] * iz + K[3]);
  // d project / d P  (calibu linear camera)
  const double a = K[0] * iz, b = K[1] * iz;
  const double Jp[6] = {a, 0.0, -a * Xs[0] * iz, 0.0, b, -b * Xs[1] * iz};
  double B[6];  // Jp * R_sw
#pragma unroll
  for (int cc = 0; cc < 3; ++cc) {
    B[cc] = Jp[0] * R[cc] + Jp[2] * R[6 + cc];
    B[3 + cc] = Jp[4] * R[3 + cc] + Jp[5] * R[6 + cc];
  }
  if (L == 3) {
#pragma unroll
    for (int i = 0; i < 6; ++i) o.Jl[i] = -B[i];
  } else {
    o.Jl[0] = -(B[0] * c.tws_r[0] + B[1] * c.tws_r[1] + B[2] * c.tws_r[2] + Jp[0] * t[0] + Jp[2] * t[2]);
    o.Jl[1] = -(B[3] * c.tws_r[0] + B[4] * c.tws_r[1] + B[5] * c.tws_r[2] + Jp[4] * t[1] + Jp[5] * t[2]);
  }
  const bool diff = (m != c.ref_pose);
  const int mo = P.pose_opt[m];
  o.pm = (diff && mo >= 0) ? mo : -1;
  o.ref_on = (L == 1) && diff && c.ref_opt >= 0;
  if (o.pm >= 0) {
    // d z / d x_meas = [ rho B | -Jp R_sv [X_v]x ],  X_v = R_vs X_s + rho t_vs
    const double* Rvs = P.cam_Rvs + 9 * cam;
    const double* tvs = P.cam_tvs + 3 * cam;
    double Xv[3];
    mat3_vec(Rvs, Xs, Xv);
    Xv[0] += c.rho * tvs[0]; Xv[1] += c.rho * tvs[1]; Xv[2] += c.rho * tvs[2];
    double C[6];  // Jp * R_vs^T
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) {
      C[cc] = Jp[0] * Rvs[3 * cc] + Jp[2] * Rvs[3 * cc + 2];
      C[3 + cc] = Jp[4] * Rvs[3 * cc + 1] + Jp[5] * Rvs[3 * cc + 2];
    }
    const uint32_t fr = P.pose_free[m];
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const double* Cr = C + 3 * rr;
      o.Jm[6 * rr + 0] = c.rho * B[3 * rr + 0];
      o.Jm[6 * rr + 1] = c.rho * B[3 * rr + 1];
      o.Jm[6 * rr + 2] = c.rho * B[3 * rr + 2];
      o.Jm[6 * rr + 3] = -(Cr[1] * Xv[2] - Cr[2] * Xv[1]);
      o.Jm[6 * rr + 4] = -(-Cr[0] * Xv[2] + Cr[2] * Xv[0]);
      o.Jm[6 * rr + 5] = -(Cr[0] * Xv[1] - Cr[1] * Xv[0]);
    }
    if ((fr & 63u) != 63u) {
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (!((fr >> k) & 1u)) { o.Jm[k] = 0.0; o.Jm[6 + k] = 0.0; }
    }
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) o.Bp[i] = B[i];
}

// reference-pose Jacobian of an observation (LmSize 1, o.ref_on): d z / d x_ref = Bp [ -rho I | R_wp_r [Y]x ], columns of
// masked parameters zeroed (linear in Bp: a scaled Bp gives the scaled Jacobian)
template <int L>
BA_D void ref_jacobian(const LmCtx<L>& c, const double* Bp, double* Jr) {
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const double* Br = Bp + 3 * rr;
    Jr[6 * rr + 0] = -c.rho * Br[0];
    Jr[6 * rr + 1] = -c.rho * Br[1];
    Jr[6 * rr + 2] = -c.rho * Br[2];
    Jr[6 * rr + 3] = Br[0] * c.Ar[0] + Br[1] * c.Ar[3] + Br[2] * c.Ar[6];
    Jr[6 * rr + 4] = Br[0] * c.Ar[1] + Br[1] * c.Ar[4] + Br[2] * c.Ar[7];
    Jr[6 * rr + 5] = Br[0] * c.Ar[2] + Br[1] * c.Ar[5] + Br[2] * c.Ar[8];
  }
  if ((c.ref_free & 63u) != 63u) {
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (!((c.ref_free >> k) & 1u)) { Jr[k] = 0.0; Jr[6 + k] = 0.0; }
  }
}

// cpp:1372-1382: e = sqrt(maha0); e > c_huber -> weight *= c_huber / e.  The square root is taken only for residuals
// beyond c_huber^2 (c2): inliers cost one comparison.
BA_D double huber_weight(double maha0, double w0, double c_huber, double c2, int use_robust) {
  if ((use_robust & 1) && maha0 > c2) {
    const double e = sqrt(maha0);
    if (e > c_huber) return w0 * (c_huber / e);
  }
  return w0;
}

// One side of a pose pair, re-evaluated from the staged landmark record `lm` (X | rho | LmSize 1: t_ws of the
// reference frame), the observation (z, orig weight) and the frame of pose m, which is the same for every lane of
// the warp (uniform, L1-resident loads).  Same closed forms as obs_jacobians.
template <int L>
struct Side {
  double Jm[12];            // d z / d x_meas, 2 x 6, masked parameter columns zeroed
  double B[6];              // d project / d X_world = Jp R_sw, 2 x 3  (LmSize 3: J_l = -B)
  double jl[2];             // LmSize 1: d z / d rho
  double r[2];
  double w;                 // Huber weight
};
template <int L>
BA_D void side_eval(const DeviceProblem& P, const double* __restrict__ lm, double zx, double zy, double w0, int m, int cam, double c_huber,
                    double c2, Side<L>& s) {
  double T[12];
  load_frame(P.Tsw + ((size_t)m * P.n_cams + cam) * 12, T);
  const double X0 = lm[0], X1 = lm[1], X2 = lm[2], rho = lm[3];
  const double Xs0 = T[0] * X0 + T[1] * X1 + T[2] * X2 + rho * T[9];
  const double Xs1 = T[3] * X0 + T[4] * X1 + T[5] * X2 + rho * T[10];
  const double Xs2 = T[6] * X0 + T[7] * X1 + T[8] * X2 + rho * T[11];
  const double2 k01 = reinterpret_cast<const double2*>(P.cam_params + 4 * cam)[0], k23 = reinterpret_cast<const double2*>(P.cam_params + 4 * cam)[1];
  const double iz = 1.0 / Xs2;
  s.r[0] = zx - (k01.x * Xs0 * iz + k23.x);
  s.r[1] = zy - (k01.y * Xs1 * iz + k23.y);
  const double a = k01.x * iz, b = k01.y * iz;
  const double j02 = -a * Xs0 * iz, j12 = -b * Xs1 * iz;
#pragma unroll
  for (int cc = 0; cc < 3; ++cc) {
    s.B[cc] = a * T[cc] + j02 * T[6 + cc];
    s.B[3 + cc] = b * T[3 + cc] + j12 * T[6 + cc];
  }
  if (L == 1) {
    s.jl[0] = -(s.B[0] * lm[4] + s.B[1] * lm[5] + s.B[2] * lm[6] + a * T[9] + j02 * T[11]);
    s.jl[1] = -(s.B[3] * lm[4] + s.B[4] * lm[5] + s.B[5] * lm[6] + b * T[10] + j12 * T[11]);
  }
  double Xv0 = Xs0, Xv1 = Xs1, Xv2 = Xs2;
  double C[6] = {a, 0.0, j02, 0.0, b, j12};      // Jp R_vs^T
  if (!P.cam_identity) {
    const double* Rvs = P.cam_Rvs + 9 * cam;
    const double* tvs = P.cam_tvs + 3 * cam;
    Xv0 = Rvs[0] * Xs0 + Rvs[1] * Xs1 + Rvs[2] * Xs2 + rho * tvs[0];
    Xv1 = Rvs[3] * Xs0 + Rvs[4] * Xs1 + Rvs[5] * Xs2 + rho * tvs[1];
    Xv2 = Rvs[6] * Xs0 + Rvs[7] * Xs1 + Rvs[8] * Xs2 + rho * tvs[2];
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) {
      C[cc] = a * Rvs[3 * cc] + j02 * Rvs[3 * cc + 2];
      C[3 + cc] = b * Rvs[3 * cc + 1] + j12 * Rvs[3 * cc + 2];
    }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const double* Cr = C + 3 * rr;
    s.Jm[6 * rr + 0] = rho * s.B[3 * rr + 0];
    s.Jm[6 * rr + 1] = rho * s.B[3 * rr + 1];
    s.Jm[6 * rr + 2] = rho * s.B[3 * rr + 2];
    s.Jm[6 * rr + 3] = -(Cr[1] * Xv2 - Cr[2] * Xv1);
    s.Jm[6 * rr + 4] = -(-Cr[0] * Xv2 + Cr[2] * Xv0);
    s.Jm[6 * rr + 5] = -(Cr[0] * Xv1 - Cr[1] * Xv0);
  }
  const uint32_t fr = P.pose_free[m];
  if ((fr & 63u) != 63u) {
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (!((fr >> k) & 1u)) { s.Jm[k] = 0.0; s.Jm[6 + k] = 0.0; }
  }
  const double n2 = s.r[0] * s.r[0] + s.r[1] * s.r[1];
  s.w = huber_weight(n2 * w0, w0, c_huber, c2, P.use_robust);
}

// M = [-rho I | R_wp_r [Y]x] (3 x 6, masked parameter columns zeroed), Y = R_vs ray + rho t_vs: the reference-pose
// Jacobian of an observation of the landmark is Bp M (proj_kernels.cuh, ref_jacobian)
BA_D void ref_M(const DeviceProblem& P, int ref_pose, int ref_cam, double rho, const double* ray, double* M) {
  double Y[3] = {ray[0], ray[1], ray[2]};
  if (!P.cam_identity) {
    mat3_vec(P.cam_Rvs + 9 * ref_cam, ray, Y);
    Y[0] += rho * P.cam_tvs[3 * ref_cam]; Y[1] += rho * P.cam_tvs[3 * ref_cam + 1]; Y[2] += rho * P.cam_tvs[3 * ref_cam + 2];
  }
  double Ar[9];
  mat3_skew(P.Rwp + 9 * (size_t)ref_pose, Y, Ar);
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int k = 0; k < 3; ++k) { M[6 * j + k] = (j == k) ? -rho : 0.0; M[6 * j + 3 + k] = Ar[3 * j + k]; }
  const uint32_t fr = P.pose_free[ref_pose];
  if ((fr & 63u) != 63u) {
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (!((fr >> k) & 1u)) { M[k] = 0.0; M[6 + k] = 0.0; M[12 + k] = 0.0; }
  }
}

// Pose block (i, j), i <= j (opt ids), inside the tiled S: element (r, c) lives at p[r * rs + c * cs].
// Tiles are stored for the upper triangle of the PERMUTED tile order; when the permutation swaps the two
// tiles the block is stored transposed.
struct SBlock { double* p; int rs, cs; };
BA_D SBlock s_block(const DeviceProblem& P, int i, int j) {
  const int a = P.tile_perm[i / kTilePoses], b = P.tile_perm[j / kTilePoses];
  const int oi = (i % kTilePoses) * P.D, oj = (j % kTilePoses) * P.D;
  SBlock o;
  if (a <= b) {
    o.p = P.S_tiles + (size_t)P.tile_map[a * P.nT + b] * P.T * P.T + (size_t)oj * P.T + oi;
    o.rs = 1; o.cs = P.T;
  } else {
    o.p = P.S_tiles + (size_t)P.tile_map[b * P.nT + a] * P.T * P.T + (size_t)oi * P.T + oj;
    o.rs = P.T; o.cs = 1;
  }
  return o;
}
// diagonal element idx (natural order) of S
BA_D double* s_diag(const DeviceProblem& P, int idx) {
  const int t = P.tile_perm[idx / P.T], o = idx % P.T;
  return P.S_tiles + (size_t)P.tile_map[t * P.nT + t] * P.T * P.T + (size_t)o * P.T + o;
}

}  // namespace bagpu
