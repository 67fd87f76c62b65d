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
// 2x4 * 4x7 * 7x7 * 7x6 chain (derivation in DESIGN.md §3); J is never stored.
//
// Work decomposition: observations are sorted by landmark; a group of G lanes (G = 4..32,
// chosen from the mean track length) owns one landmark, lanes stride over its observations,
// V / b_l (and the reference-pose blocks for LmSize 1) are reduced with warp shuffles, and the
// pose-block contributions go to S / b_p with FP64 reductions (RED.E.ADD.F64) into L2.
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

BA_D double huber_weight(double maha0, double w0, double c_huber, int use_robust) {
  const double e = sqrt(maha0);
  return ((use_robust & 1) && e > c_huber) ? w0 * (c_huber / e) : w0;
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
