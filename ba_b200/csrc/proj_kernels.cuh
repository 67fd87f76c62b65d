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

// 256-bit read-only load (32-byte aligned).  The gathers of per-pose data are scattered over the lanes of a warp, and L1
// looks up one cache line per cycle whatever the access width: the fewer, wider instructions a gather takes, the better.
BA_D void ld4(const double* p, double& a, double& b, double& c, double& d) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}
// a pose frame [R (9, row-major) | t (3)] in three 32-byte loads (frames are 96 bytes apart)
BA_D void load_frame(const double* __restrict__ T, double* f) {
  ld4(T, f[0], f[1], f[2], f[3]); ld4(T + 4, f[4], f[5], f[6], f[7]); ld4(T + 8, f[8], f[9], f[10], f[11]);
}

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
    double Tws[12];
    load_frame(P.Tws + ((size_t)c.ref_pose * P.n_cams + rc) * 12, Tws);
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

// Linearisation of one observation for the Schur build (ba_gpu.h): B = d project / d X_world = Jp R_sw, the Huber
// weight, G = w B^T B, h = w B^T r and, LmSize 1, j = B u = -d z / d rho (u = baseline reference sensor -> measuring
// sensor; B u = B t_ws_ref + Jp t_sw without forming u).
template <int L>
struct ObsLin {
  double G[6];      // upper: 00 01 02 11 12 22
  double h[3];
  double j[2];      // LmSize 1: B u
  double gj[3];     // LmSize 1: G u = w B^T j
  double wjr;       // LmSize 1: u^T h = w j.r
  double w, n2;     // Huber weight, |r|^2
  bool pm_on;       // measurement pose takes part (optimised, not the landmark's reference pose)
  bool ref_on;      // reference pose takes part (LmSize 1, optimised, diff_poses)
};
template <int L>
BA_D void obs_linearize(const DeviceProblem& P, const LmCtx<L>& c, int obs, int m, double c_huber, double c2, ObsLin<L>& o) {
  const int cam = P.obs_cam ? (int)P.obs_cam[obs] : 0;
  double T[12];
  load_frame(P.Tsw + ((size_t)m * P.n_cams + cam) * 12, T);
  const double Xs0 = T[0] * c.X[0] + T[1] * c.X[1] + T[2] * c.X[2] + c.rho * T[9];
  const double Xs1 = T[3] * c.X[0] + T[4] * c.X[1] + T[5] * c.X[2] + c.rho * T[10];
  const double Xs2 = T[6] * c.X[0] + T[7] * c.X[1] + T[8] * c.X[2] + c.rho * T[11];
  const double2 k01 = reinterpret_cast<const double2*>(P.cam_params + 4 * cam)[0], k23 = reinterpret_cast<const double2*>(P.cam_params + 4 * cam)[1];
  const double iz = 1.0 / Xs2;
  const double2 z = P.obs_z[obs];
  const double r0 = z.x - (k01.x * Xs0 * iz + k23.x), r1 = z.y - (k01.y * Xs1 * iz + k23.y);
  const double a = k01.x * iz, b = k01.y * iz;
  const double j02 = -a * Xs0 * iz, j12 = -b * Xs1 * iz;
  double B[6];
#pragma unroll
  for (int cc = 0; cc < 3; ++cc) { B[cc] = a * T[cc] + j02 * T[6 + cc]; B[3 + cc] = b * T[3 + cc] + j12 * T[6 + cc]; }
  if (L == 1) {
    o.j[0] = B[0] * c.tws_r[0] + B[1] * c.tws_r[1] + B[2] * c.tws_r[2] + a * T[9] + j02 * T[11];
    o.j[1] = B[3] * c.tws_r[0] + B[4] * c.tws_r[1] + B[5] * c.tws_r[2] + b * T[10] + j12 * T[11];
  }
  const double w0 = P.obs_w0[obs];
  o.n2 = r0 * r0 + r1 * r1;
  o.w = huber_weight(o.n2 * w0, w0, c_huber, c2, P.use_robust);
  const double wb0 = o.w * B[0], wb1 = o.w * B[1], wb2 = o.w * B[2], wb3 = o.w * B[3], wb4 = o.w * B[4], wb5 = o.w * B[5];
  o.G[0] = wb0 * B[0] + wb3 * B[3]; o.G[1] = wb0 * B[1] + wb3 * B[4]; o.G[2] = wb0 * B[2] + wb3 * B[5];
  o.G[3] = wb1 * B[1] + wb4 * B[4]; o.G[4] = wb1 * B[2] + wb4 * B[5]; o.G[5] = wb2 * B[2] + wb5 * B[5];
  o.h[0] = wb0 * r0 + wb3 * r1; o.h[1] = wb1 * r0 + wb4 * r1; o.h[2] = wb2 * r0 + wb5 * r1;
  if (L == 1) {
    o.gj[0] = wb0 * o.j[0] + wb3 * o.j[1]; o.gj[1] = wb1 * o.j[0] + wb4 * o.j[1]; o.gj[2] = wb2 * o.j[0] + wb5 * o.j[1];
    o.wjr = o.w * (o.j[0] * r0 + o.j[1] * r1);
  }
  const bool diff = (m != c.ref_pose);
  o.pm_on = diff && P.pose_opt[m] >= 0;
  o.ref_on = (L == 1) && diff && c.ref_opt >= 0;
}

// ---- K~^T M K~ products: K~(Y) = [rho I | -[Y]x] (3 x 6), Y = X - rho t_wp ----------------------------------------
// symmetric 3 x 3 (upper 00 01 02 11 12 22) times vector
BA_D void sym3_vec(const double* G, const double* v, double* o) {
  o[0] = G[0] * v[0] + G[1] * v[1] + G[2] * v[2];
  o[1] = G[1] * v[0] + G[3] * v[1] + G[4] * v[2];
  o[2] = G[2] * v[0] + G[4] * v[1] + G[5] * v[2];
}
// p = K~^T g = [rho g ; Y x g]
BA_D void kt_vec(double rho, const double* Y, const double* g, double* p) {
  p[0] = rho * g[0]; p[1] = rho * g[1]; p[2] = rho * g[2];
  p[3] = Y[1] * g[2] - Y[2] * g[1]; p[4] = Y[2] * g[0] - Y[0] * g[2]; p[5] = Y[0] * g[1] - Y[1] * g[0];
}
// N = M K~(Yb)  (3 x 6, row-major): [rho M | -M [Yb]x]
BA_D void m_times_k(const double* M, double rho, const double* Yb, double* N) {
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double m0 = M[3 * r], m1 = M[3 * r + 1], m2 = M[3 * r + 2];
    N[6 * r] = rho * m0; N[6 * r + 1] = rho * m1; N[6 * r + 2] = rho * m2;
    N[6 * r + 3] = m2 * Yb[1] - m1 * Yb[2];
    N[6 * r + 4] = m0 * Yb[2] - m2 * Yb[0];
    N[6 * r + 5] = m1 * Yb[0] - m0 * Yb[1];
  }
}
// acc (6 x 6 row-major) += K~(Ya)^T M K~(Yb), M full 3 x 3 row-major.  Column by column: column c of M K~(Yb) is
// rho M[:, c] (c < 3) or M (e_{c-3} x Yb), three values live at a time.
BA_D void acc_kmk(double* acc, double rho, const double* Ya, const double* Yb, const double* M) {
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    double n0, n1, n2;
    if (c < 3) { n0 = rho * M[c]; n1 = rho * M[3 + c]; n2 = rho * M[6 + c]; }
    else {
      const int p = (c - 3 + 2) % 3, q = (c - 3 + 1) % 3;   // column = M[:, p] Yb[q] - M[:, q] Yb[p]
      n0 = M[p] * Yb[q] - M[q] * Yb[p]; n1 = M[3 + p] * Yb[q] - M[3 + q] * Yb[p]; n2 = M[6 + p] * Yb[q] - M[6 + q] * Yb[p];
    }
    acc[c] += rho * n0; acc[6 + c] += rho * n1; acc[12 + c] += rho * n2;
    acc[18 + c] += Ya[1] * n2 - Ya[2] * n1;
    acc[24 + c] += Ya[2] * n0 - Ya[0] * n2;
    acc[30 + c] += Ya[0] * n1 - Ya[1] * n0;
  }
}
// acc (upper triangle packed k = c (c + 1) / 2 + r, r <= c) += K~(Y)^T M K~(Y), M symmetric given as full row-major
BA_D void acc_kmk_sym(double* acc, double rho, const double* Y, const double* M) {
  double N[18];
  m_times_k(M, rho, Y, N);
  int k = 0;
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const double n0 = N[c], n1 = N[6 + c], n2 = N[12 + c];
    const double col[6] = {rho * n0, rho * n1, rho * n2, Y[1] * n2 - Y[2] * n1, Y[2] * n0 - Y[0] * n2, Y[0] * n1 - Y[1] * n0};
#pragma unroll
    for (int r = 0; r <= c; ++r) { acc[k] += col[r]; ++k; }
  }
}
// y = K_pose delta = K~ (D delta): x_t = mask . delta[0:3], x_w = R_wp (mask . delta[3:6]), y = rho x_t - (X - rho t_wp) x x_w
BA_D void k_times_delta(const DeviceProblem& P, int pose, const double* X, double rho, const double* __restrict__ delta, double* y) {
  const uint32_t fr = P.pose_free[pose];
  double x[6];
#pragma unroll
  for (int a = 0; a < 6; ++a) x[a] = ((fr >> a) & 1u) ? delta[a] : 0.0;
  double xw[3];
  mat3_vec(P.Rwp + 9 * (size_t)pose, x + 3, xw);
  const double* tp = P.pose_t + 3 * (size_t)pose;
  const double Y[3] = {X[0] - rho * tp[0], X[1] - rho * tp[1], X[2] - rho * tp[2]};
  y[0] = rho * x[0] - (Y[1] * xw[2] - Y[2] * xw[1]);
  y[1] = rho * x[1] - (Y[2] * xw[0] - Y[0] * xw[2]);
  y[2] = rho * x[2] - (Y[0] * xw[1] - Y[1] * xw[0]);
}
// The same product from the pose's PACK (k_pose_delta_pack: 12 doubles = three 32-byte loads per pose instead of ~20 scalar
// gathers): [x_t (3) | x_w (3) | t_wp (3) | t_ws of camera 0 (3)], x_t[0] = NaN for a pose that is not optimised.
// Same arithmetic in the same order as k_times_delta.  Returns false (y = 0) for a pose that is not optimised.
constexpr int kPosePack = 12;
BA_D bool k_times_pack(const double* __restrict__ pk, const double* X, double rho, double* y, double* tw) {
  double x[12];
  load_frame(pk, x);
  tw[0] = x[9]; tw[1] = x[10]; tw[2] = x[11];
  if (x[0] != x[0]) { y[0] = 0.0; y[1] = 0.0; y[2] = 0.0; return false; }
  const double Y[3] = {X[0] - rho * x[6], X[1] - rho * x[7], X[2] - rho * x[8]};
  y[0] = rho * x[0] - (Y[1] * x[5] - Y[2] * x[4]);
  y[1] = rho * x[1] - (Y[2] * x[3] - Y[0] * x[5]);
  y[2] = rho * x[2] - (Y[0] * x[4] - Y[1] * x[3]);
  return true;
}
// what the kernels behind the build need of a landmark, from the record the linearisation wrote (state unchanged since)
struct LmRecCtx { double X[3], rho, tws_r[3]; int ref_pose; };
template <int L, bool WITH_TWS = true>
BA_D void load_lm_rec(const DeviceProblem& P, int lm, LmRecCtx& c) {
  const double* r = P.lm_rec + (size_t)lm * kLmRec;
  ld4(r, c.X[0], c.X[1], c.X[2], c.rho);
  if (L == 1 && WITH_TWS) { double vi; ld4(r + 4, c.tws_r[0], c.tws_r[1], c.tws_r[2], vi); (void)vi; }
  else { c.tws_r[0] = c.tws_r[1] = c.tws_r[2] = 0.0; }
  c.ref_pose = (int)P.lm_ref_pose[lm];
}
// the G record of an observation (upper 00 01 02 11 12 22): three 16-byte loads
BA_D void load_G(const DeviceProblem& P, int obs, double* G) {
  const double2* g = reinterpret_cast<const double2*>(P.obs_G + (size_t)obs * kObsG);
  const double2 a = g[0], b = g[1], c = g[2];
  G[0] = a.x; G[1] = a.y; G[2] = b.x; G[3] = b.y; G[4] = c.x; G[5] = c.y;
}

// D_p = blkdiag(I, R_wp) with masked parameter columns zeroed: element (q, c), R row-major
BA_D double d_elem(const double* R, uint32_t free_bits, int q, int c) {
  if (!((free_bits >> c) & 1u)) return 0.0;
  if (c < 3) return q == c ? 1.0 : 0.0;
  return q >= 3 ? R[3 * (q - 3) + (c - 3)] : 0.0;
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
