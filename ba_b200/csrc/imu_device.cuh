// imu_device.cuh — inertial residual: RK4 pre-integration with bias Jacobian and covariance,
// residual, Jacobians w.r.t. both poses, information matrix.  One thread per residual (the
// chain over the IMU samples is sequential; there are only ~P residuals).
//
// Follows the reference's definitions, including its approximations (the Jacobians of the
// integrator are the reference's own chain, e.g. dq_exp_dw, not exact derivatives):
//   IntegratePose       reference include/ba/Types.h:324-373
//   GetPoseDerivative   reference include/ba/Types.h:376-416
//   IntegrateImu        reference include/ba/Types.h:419-643   (euler_covariance = true)
//   IntegrateResidual   reference include/ba/Types.h:662-738
//   residual/Jacobians  reference include/ba/parallel_algos.h:178-358
//   weighting           reference src/BundleAdjuster.cpp:1497-1541 (quirk Q3: the Huber threshold
//                       comes from the *projection* errors; with no projection residuals the
//                       inertial block contributes nothing)
// The pose-block Jacobians dLog_decoupled_dt1 * dt1_t2_dt1 * dexp_decoupled_dx etc. are
// evaluated in closed form with the inverse right Jacobian of SO3.
#pragma once
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"
#include "proj_kernels.cuh"

namespace bagpu {

// ---- tiny dense helpers on row-major arrays ---------------------------------------------------
template <int R, int K, int C>
BA_D void mm(const double* A, const double* B, double* O) {   // O(RxC) = A(RxK) B(KxC)
  for (int r = 0; r < R; ++r)
    for (int c = 0; c < C; ++c) {
      double s = 0.0;
      for (int k = 0; k < K; ++k) s += A[r * K + k] * B[k * C + c];
      O[r * C + c] = s;
    }
}
template <int N>
BA_D void mzero(double* A) { for (int i = 0; i < N; ++i) A[i] = 0.0; }
template <int N>
BA_D void mcopy(const double* A, double* O) { for (int i = 0; i < N; ++i) O[i] = A[i]; }

// Utils.h:294-312  d(R(q) v)/dq with independent components, 3x4
BA_D void dqx_dq(const double* q, const double* v, double* m) {
  const double x = v[0], y = v[1], z = v[2];
  const double qx = q[0], qy = q[1], qz = q[2], qw = q[3];
  const double s1 = 2 * qx * y, s2 = 2 * qy * y, s3 = 2 * qx * x, s4 = 2 * qz * x, s5 = 2 * qy * z, s6 = 2 * qz * z;
  m[0] = s2 + s6; m[1] = s1 - 4 * qy * x + 2 * qw * z; m[2] = 2 * qx * z - 2 * qw * y - 4 * qz * x; m[3] = s5 - 2 * qz * y;
  m[4] = 2 * qy * x - 4 * qx * y - 2 * qw * z; m[5] = s3 + s6; m[6] = s5 + 2 * qw * x - 4 * qz * y; m[7] = s4 - 2 * qx * z;
  m[8] = s4 + 2 * qw * y - 4 * qx * z; m[9] = 2 * qz * y - 2 * qw * x - 4 * qy * z; m[10] = s2 + s3; m[11] = s1 - 2 * qy * x;
}
// Utils.h:276-291
BA_D void dq1q2_dq2(const double* q1, double* m) {
  const double x = q1[0], y = q1[1], z = q1[2], w = q1[3];
  m[0] = w; m[1] = -z; m[2] = y; m[3] = x; m[4] = z; m[5] = w; m[6] = -x; m[7] = y;
  m[8] = -y; m[9] = x; m[10] = w; m[11] = z; m[12] = -x; m[13] = -y; m[14] = -z; m[15] = w;
}
BA_D void dq1q2_dq1(const double* q2, double* m) {
  const double x = q2[0], y = q2[1], z = q2[2], w = q2[3];
  m[0] = w; m[1] = z; m[2] = -y; m[3] = x; m[4] = -z; m[5] = w; m[6] = x; m[7] = y;
  m[8] = y; m[9] = -x; m[10] = w; m[11] = z; m[12] = -x; m[13] = -y; m[14] = -z; m[15] = w;
}
// Utils.h:251-266 (4x3)
BA_D void dq_exp_dw(const double* w, double* m) {
  const double t = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  const double s1 = t / 20 - 1;
  const double s2 = t * t / 48 - 0.5;
  const double s3 = (s1 * w[1] * w[2]) / 24, s4 = (s1 * w[0] * w[2]) / 24, s5 = (s1 * w[0] * w[1]) / 24;
  const double s6 = t * t;
  m[0] = (s1 * w[0] * w[0]) / 24 - s6 / 48 + 0.5; m[1] = s5; m[2] = s4;
  m[3] = s5; m[4] = (s1 * w[1] * w[1]) / 24 - s6 / 48 + 0.5; m[5] = s3;
  m[6] = s4; m[7] = s3; m[8] = (s1 * w[2] * w[2]) / 24 - s6 / 48 + 0.5;
  m[9] = (s2 * w[0]) / 2; m[10] = (s2 * w[1]) / 2; m[11] = (s2 * w[2]) / 2;
}
// Utils.h:136-185 (3x4)
BA_D void dlog_dq(const double* q, double* m) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double n2 = x * x + y * y + z * z;
  const double n = sqrt(n2);
  if (n < 1e-9) {
    const double s1 = 2 * n2, s2 = 1.0 / (w * w * w), s3 = (3 * s1) / (w * w * w * w) - 2 / (w * w), s4 = 2 / w;
    m[0] = -4 * s2 * x * x + s4 - s1 * s2; m[1] = -4 * x * y * s2; m[2] = -4 * x * z * s2; m[3] = x * s3;
    m[4] = -4 * x * y * s2; m[5] = -4 * s2 * y * y + s4 - s1 * s2; m[6] = -4 * y * z * s2; m[7] = y * s3;
    m[8] = -4 * x * z * s2; m[9] = -4 * y * z * s2; m[10] = -4 * s2 * z * z + s4 - s1 * s2; m[11] = z * s3;
  } else {
    const double s1 = n2, s2 = 1 / (s1 / (w * w) + 1), s3 = atan(n / w), s4 = 1 / (s1 * n), s5 = 1 / s1, s6 = 1 / w;
    const double s7 = (2 * s3) / n;
    const double s8 = 2 * y * z * s2 * s5 * s6 - 2 * y * z * s3 * s4;
    const double s9 = 2 * x * z * s2 * s5 * s6 - 2 * x * z * s3 * s4;
    const double s10 = 2 * x * y * s2 * s5 * s6 - 2 * x * y * s3 * s4;
    m[0] = s7 - 2 * x * x * s3 * s4 + 2 * x * x * s2 * s5 * s6; m[1] = s10; m[2] = s9; m[3] = -(2 * x * s2) / (w * w);
    m[4] = s10; m[5] = s7 - 2 * y * y * s3 * s4 + 2 * y * y * s2 * s5 * s6; m[6] = s8; m[7] = -(2 * y * s2) / (w * w);
    m[8] = s9; m[9] = s8; m[10] = s7 - 2 * z * z * s3 * s4 + 2 * z * z * s2 * s5 * s6; m[11] = -(2 * z * s2) / (w * w);
  }
}

struct ImuState { double t[3], q[4], v[3]; };
struct ImuSample { double w[3], a[3], time; };

BA_D void load_sample(const double* m7, ImuSample& s) {
  s.w[0] = m7[0]; s.w[1] = m7[1]; s.w[2] = m7[2]; s.a[0] = m7[3]; s.a[1] = m7[4]; s.a[2] = m7[5]; s.time = m7[6];
}

// Types.h:324-373.  dy_dk: 10x9, dy_dy: 4x4 (only the quaternion block), both optional
BA_D void integrate_pose(const ImuState& p, const double* k, double dt, ImuState& y, double* dy_dk, double* dy_dy) {
  const double wdt[3] = {k[3] * dt, k[4] * dt, k[5] * dt};
  double r[4];
  so3_exp(wdt, r);
  y.t[0] = p.t[0] + k[0] * dt; y.t[1] = p.t[1] + k[1] * dt; y.t[2] = p.t[2] + k[2] * dt;
  quat_mul(r, p.q, y.q);   // not re-normalised (Types.h:336-339)
  y.v[0] = p.v[0] + k[6] * dt; y.v[1] = p.v[1] + k[7] * dt; y.v[2] = p.v[2] + k[8] * dt;
  if (dy_dk) {
    mzero<90>(dy_dk);
    dy_dk[0 * 9 + 0] = dt; dy_dk[1 * 9 + 1] = dt; dy_dk[2 * 9 + 2] = dt;
    double a[16], b[12], c[12];
    dq1q2_dq1(p.q, a);
    dq_exp_dw(wdt, b);
    mm<4, 4, 3>(a, b, c);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 3; ++j) dy_dk[(3 + i) * 9 + 3 + j] = c[i * 3 + j] * dt;
    dy_dk[7 * 9 + 6] = dt; dy_dk[8 * 9 + 7] = dt; dy_dk[9 * 9 + 8] = dt;
  }
  if (dy_dy) dq1q2_dq2(r, dy_dy);
}

// Types.h:376-416.  k: 9, dk_db: 9x6, dk_dx: 9x10
BA_D void pose_derivative(const ImuState& p, const double* g, const ImuSample& z0, const ImuSample& z1, const double* bg,
                          const double* ba, double dt, double* k, double* dk_db, double* dk_dx) {
  const double alpha = (z1.time - (z0.time + dt)) / (z1.time - z0.time);
  double zg[3], za[3];
  for (int i = 0; i < 3; ++i) { zg[i] = z0.w[i] * alpha + z1.w[i] * (1.0 - alpha); za[i] = z0.a[i] * alpha + z1.a[i] * (1.0 - alpha); }
  double R[9];
  quat_to_R(p.q, R);
  const double wg[3] = {zg[0] + bg[0], zg[1] + bg[1], zg[2] + bg[2]};
  const double aa[3] = {za[0] + ba[0], za[1] + ba[1], za[2] + ba[2]};
  double w[3], a[3];
  mat3_vec(R, wg, w);
  mat3_vec(R, aa, a);
  k[0] = p.v[0]; k[1] = p.v[1]; k[2] = p.v[2];
  k[3] = w[0]; k[4] = w[1]; k[5] = w[2];
  k[6] = a[0] - g[0]; k[7] = a[1] - g[1]; k[8] = a[2] - g[2];
  if (dk_db) {
    mzero<54>(dk_db);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { dk_db[(3 + i) * 6 + j] = R[3 * i + j]; dk_db[(6 + i) * 6 + 3 + j] = R[3 * i + j]; }
  }
  if (dk_dx) {
    mzero<90>(dk_dx);
    dk_dx[0 * 10 + 7] = 1.0; dk_dx[1 * 10 + 8] = 1.0; dk_dx[2 * 10 + 9] = 1.0;
    double m1[12], m2[12];
    dqx_dq(p.q, zg, m1); dqx_dq(p.q, bg, m2);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 4; ++j) dk_dx[(3 + i) * 10 + 3 + j] = m1[i * 4 + j] + m2[i * 4 + j];
    dqx_dq(p.q, za, m1); dqx_dq(p.q, ba, m2);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 4; ++j) dk_dx[(6 + i) * 10 + 3 + j] = m1[i * 4 + j] + m2[i * 4 + j];
  }
}

BA_D void add_identity_blocks(double* dy_dy0, const double* dy_dy) {
  for (int i = 0; i < 3; ++i) { dy_dy0[i * 10 + i] += 1.0; dy_dy0[(7 + i) * 10 + 7 + i] += 1.0; }
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) dy_dy0[(3 + i) * 10 + 3 + j] += dy_dy[i * 4 + j];
}

// Types.h:419-643 with Jacobians: returns the new state, dy_db (10x6), dy_dy0 (10x10)
BA_D void integrate_imu_jac(const ImuState& pose, const ImuSample& z0, const ImuSample& z1, const double* bg, const double* ba,
                            const double* g, ImuState& res, double* dy_db, double* dy_dy0) {
  const double dt = z1.time - z0.time;
  double dk_db[54], dk_dy[90], dy_dk[90], dy_dy[16];
  double k1[9], k2[9], k3[9], k4[9];
  double dk1_db[54], dk1_dy[90], dkt_db[54], dkt_dy[90], tmp6[54], tmp10[90];
  ImuState y;
  pose_derivative(pose, g, z0, z1, bg, ba, 0.0, k1, dk_db, dk_dy);
  mcopy<54>(dk_db, dk1_db); mcopy<90>(dk_dy, dk1_dy);
  mcopy<54>(dk1_db, dkt_db); mcopy<90>(dk1_dy, dkt_dy);            // running total: k1 + 2 k2 + 2 k3 + k4
  integrate_pose(pose, k1, dt * 0.5, y, dy_dk, dy_dy);
  mm<10, 9, 6>(dy_dk, dk1_db, dy_db);
  mm<10, 9, 10>(dy_dk, dk1_dy, dy_dy0);
  add_identity_blocks(dy_dy0, dy_dy);
  // k2
  pose_derivative(y, g, z0, z1, bg, ba, dt / 2, k2, dk_db, dk_dy);
  mm<9, 10, 6>(dk_dy, dy_db, tmp6);
  for (int i = 0; i < 54; ++i) { tmp6[i] += dk_db[i]; dkt_db[i] += 2 * tmp6[i]; }      // dk2_db
  mm<9, 10, 10>(dk_dy, dy_dy0, tmp10);
  for (int i = 0; i < 90; ++i) dkt_dy[i] += 2 * tmp10[i];                              // dk2_dy
  integrate_pose(pose, k2, dt * 0.5, y, dy_dk, dy_dy);
  mm<10, 9, 6>(dy_dk, tmp6, dy_db);
  mm<10, 9, 10>(dy_dk, tmp10, dy_dy0);
  add_identity_blocks(dy_dy0, dy_dy);
  // k3
  pose_derivative(y, g, z0, z1, bg, ba, dt / 2, k3, dk_db, dk_dy);
  mm<9, 10, 6>(dk_dy, dy_db, tmp6);
  for (int i = 0; i < 54; ++i) { tmp6[i] += dk_db[i]; dkt_db[i] += 2 * tmp6[i]; }
  mm<9, 10, 10>(dk_dy, dy_dy0, tmp10);
  for (int i = 0; i < 90; ++i) dkt_dy[i] += 2 * tmp10[i];
  integrate_pose(pose, k3, dt, y, dy_dk, dy_dy);
  mm<10, 9, 6>(dy_dk, tmp6, dy_db);
  mm<10, 9, 10>(dy_dk, tmp10, dy_dy0);
  add_identity_blocks(dy_dy0, dy_dy);
  // k4
  pose_derivative(y, g, z0, z1, bg, ba, dt, k4, dk_db, dk_dy);
  mm<9, 10, 6>(dk_dy, dy_db, tmp6);
  for (int i = 0; i < 54; ++i) dkt_db[i] += tmp6[i] + dk_db[i];
  mm<9, 10, 10>(dk_dy, dy_dy0, tmp10);
  for (int i = 0; i < 90; ++i) dkt_dy[i] += tmp10[i];
  double k[9];
  for (int i = 0; i < 9; ++i) k[i] = k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i];
  integrate_pose(pose, k, dt / 6.0, res, dy_dk, dy_dy);
  mm<10, 9, 6>(dy_dk, dkt_db, dy_db);
  mm<10, 9, 10>(dy_dk, dkt_dy, dy_dy0);
  add_identity_blocks(dy_dy0, dy_dy);
}

// Types.h:618-637 (no Jacobians)
BA_D void integrate_imu(const ImuState& pose, const ImuSample& z0, const ImuSample& z1, const double* bg, const double* ba,
                        const double* g, ImuState& res) {
  const double dt = z1.time - z0.time;
  double k1[9], k2[9], k3[9], k4[9], k[9];
  ImuState y;
  pose_derivative(pose, g, z0, z1, bg, ba, 0.0, k1, nullptr, nullptr);
  integrate_pose(pose, k1, dt * 0.5, y, nullptr, nullptr);
  pose_derivative(y, g, z0, z1, bg, ba, dt / 2, k2, nullptr, nullptr);
  integrate_pose(pose, k2, dt * 0.5, y, nullptr, nullptr);
  pose_derivative(y, g, z0, z1, bg, ba, dt / 2, k3, nullptr, nullptr);
  integrate_pose(pose, k3, dt, y, nullptr, nullptr);
  pose_derivative(y, g, z0, z1, bg, ba, dt, k4, nullptr, nullptr);
  for (int i = 0; i < 9; ++i) k[i] = k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i];
  integrate_pose(pose, k, dt / 6.0, res, nullptr, nullptr);
}

BA_D void load_state(const DeviceProblem& P, int p, ImuState& s, double* b) {
  for (int k = 0; k < 4; ++k) s.q[k] = P.pose_q[4 * (size_t)p + k];
  for (int k = 0; k < 3; ++k) { s.t[k] = P.pose_t[3 * (size_t)p + k]; s.v[k] = P.pose_v[3 * (size_t)p + k]; }
  for (int k = 0; k < 6; ++k) b[k] = P.pose_b[6 * (size_t)p + k];
}

// residual (D) of the inertial constraint given the integrated state (parallel_algos.h:267-270,336)
template <int D>
BA_D void imu_residual(const ImuState& yi, const ImuState& s2, const double* b1, const double* b2, double* r, double* phi, double* qQ) {
  const double q2c[4] = {-s2.q[0], -s2.q[1], -s2.q[2], s2.q[3]};
  quat_mul(yi.q, q2c, qQ); quat_normalize(qQ);
  so3_log(qQ, phi);
  for (int k = 0; k < D; ++k) r[k] = 0.0;
  for (int k = 0; k < 3; ++k) { r[k] = yi.t[k] - s2.t[k]; r[3 + k] = phi[k]; r[6 + k] = yi.v[k] - s2.v[k]; }
  if (D >= 15) for (int k = 0; k < 6; ++k) r[9 + k] = b1[k] - b2[k];
}

// in-place inverse by Gauss-Jordan with partial pivoting (stands in for Eigen's inverse(), parallel_algos.h:307)
template <int N>
BA_D void invert(double* M, double* I) {
  for (int i = 0; i < N * N; ++i) I[i] = 0.0;
  for (int i = 0; i < N; ++i) I[i * N + i] = 1.0;
  for (int c = 0; c < N; ++c) {
    int p = c; double best = fabs(M[c * N + c]);
    for (int r = c + 1; r < N; ++r) if (fabs(M[r * N + c]) > best) { best = fabs(M[r * N + c]); p = r; }
    if (p != c) for (int j = 0; j < N; ++j) { double t = M[c * N + j]; M[c * N + j] = M[p * N + j]; M[p * N + j] = t; t = I[c * N + j]; I[c * N + j] = I[p * N + j]; I[p * N + j] = t; }
    const double d = 1.0 / M[c * N + c];
    for (int j = 0; j < N; ++j) { M[c * N + j] *= d; I[c * N + j] *= d; }
    for (int r = 0; r < N; ++r) if (r != c) {
      const double f = M[r * N + c];
      if (f != 0.0) for (int j = 0; j < N; ++j) { M[r * N + j] -= f * M[c * N + j]; I[r * N + j] -= f * I[c * N + j]; }
    }
  }
}

template <int D>
BA_D double imu_build_one(const DeviceProblem& P, const OtherResiduals& R, int i) {   // returns this residual's share of inertial_error
  const int p1 = (int)R.imu_p1[i], p2 = (int)R.imu_p2[i];
  ImuState s1, s2, y;
  double b1[6], b2[6];
  load_state(P, p1, s1, b1); load_state(P, p2, s2, b2);
  const double* g = R.gravity;
  y = s1;
  double db[60], C[100];
  mzero<60>(db); mzero<100>(C);
  const int m0 = (int)R.imu_meas_start[i], m1 = (int)R.imu_meas_start[i + 1];
  ImuSample za, zb;
  for (int m = m0 + 1; m < m1; ++m) {
    load_sample(R.imu_meas + 7 * (size_t)(m - 1), za); load_sample(R.imu_meas + 7 * (size_t)m, zb);
    if (zb.time - za.time == 0) continue;
    ImuState yn;
    double dy_db[60], dy_dy0[100], t1[100], t2[100];
    integrate_imu_jac(y, za, zb, b1, b1 + 3, g, yn, dy_db, dy_dy0);
    y = yn;
    // c = dy_dy0 c dy_dy0^T + dy_db r dy_db^T   (Types.h:601-606)
    mm<10, 10, 10>(dy_dy0, C, t1);
    for (int r = 0; r < 10; ++r) for (int c = 0; c < 10; ++c) { double s = 0.0; for (int k = 0; k < 10; ++k) s += t1[r * 10 + k] * dy_dy0[c * 10 + k]; t2[r * 10 + c] = s; }
    for (int r = 0; r < 10; ++r) for (int c = 0; c < 10; ++c) { double s = 0.0; for (int k = 0; k < 6; ++k) s += dy_db[r * 6 + k] * R.imu_r[k] * dy_db[c * 6 + k]; C[r * 10 + c] = t2[r * 10 + c] + s; }
    // dintegration_db = dy_db + dy_dy0 * dintegration_db   (Types.h:712-714)
    double t3[60];
    mm<10, 10, 6>(dy_dy0, db, t3);
    for (int k = 0; k < 60; ++k) db[k] = dy_db[k] + t3[k];
  }
  if (R.cov_once) {   // res.covariance_computed: later builds keep the first build's dintegration_db / c_integration
    double* keep = R.imu_once + (size_t)i * 160;
    if (R.imu_once_done[i]) { for (int k = 0; k < 60; ++k) db[k] = keep[k]; for (int k = 0; k < 100; ++k) C[k] = keep[60 + k]; }
    else { for (int k = 0; k < 60; ++k) keep[k] = db[k]; for (int k = 0; k < 100; ++k) keep[60 + k] = C[k]; R.imu_once_done[i] = 1; }
  }
  ImuSample zf, zl;
  load_sample(R.imu_meas + 7 * (size_t)m0, zf); load_sample(R.imu_meas + 7 * (size_t)(m1 - 1), zl);
  const double total_dt = zl.time - zf.time;
  double R1[9], R2[9];
  quat_to_R(s1.q, R1); quat_to_R(s2.q, R2);
  // t_12_0 and v_12_0 (parallel_algos.h:217-227)
  double tt[3], t0[3], v0w[3], v0[3];
  for (int k = 0; k < 3; ++k) tt[k] = y.t[k] - (-g[k] * 0.5 * total_dt * total_dt + s1.v[k] * total_dt) - s1.t[k];
  mat3T_vec(R1, tt, t0);
  for (int k = 0; k < 3; ++k) v0w[k] = y.v[k] - s1.v[k] + g[k] * total_dt;
  mat3T_vec(R1, v0w, v0);
  double r[D], phi[3], qQ[4];
  imu_residual<D>(y, s2, b1, b2, r, phi, qQ);
  double J1[D * D], J2[D * D];
  mzero<D * D>(J1); mzero<D * D>(J2);
  // rotation blocks: d log(R_int R2^T)/d w1 = Jr^-1(phi) R2 R0^T, R0 = R1^T R_int ; d/d w2 = -Jr^-1(phi) R2
  double Jr[9], Ri[9], R0[9], A[9], Bm[9];
  so3_jinv(phi, 0.5, Jr);
  { double qn[4] = {y.q[0], y.q[1], y.q[2], y.q[3]}; quat_normalize(qn); quat_to_R(qn, Ri); }
  mat3_mulT(R1, Ri, R0);               // R0 = R1^T Ri
  mat3_mul(Jr, R2, A);                 // A = Jr^-1 R2
  for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 3; ++c) { double s = 0.0; for (int k = 0; k < 3; ++k) s += A[3 * rr + k] * R0[3 * c + k]; Bm[3 * rr + c] = s; }   // A R0^T
  double R1t0x[9];
  mat3_skew(R1, t0, R1t0x);            // R1 [t0]x
  for (int rr = 0; rr < 3; ++rr) {
    J1[rr * D + rr] = 1.0;
    for (int c = 0; c < 3; ++c) {
      J1[rr * D + 3 + c] = -R1t0x[3 * rr + c];
      J1[(3 + rr) * D + 3 + c] = Bm[3 * rr + c];
      J2[(3 + rr) * D + 3 + c] = -A[3 * rr + c];
    }
    J2[rr * D + rr] = -1.0;
    J1[rr * D + 6 + rr] = total_dt;       // parallel_algos.h:238-239
    J1[(6 + rr) * D + 6 + rr] = 1.0;      // :247-248
    J2[(6 + rr) * D + 6 + rr] = -1.0;     // :264-265
  }
  double R1v0x[9];
  mat3_skew(R1, v0, R1v0x);            // column k of (R1 hat(e_k) v0) = -R1 [v0]x e_k   (:240-244)
  for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 3; ++c) J1[(6 + rr) * D + 3 + c] = -R1v0x[3 * rr + c];
  // covariance of the residual: [dlog_dt1 0; 0 I] c_integration [..]^T  (parallel_algos.h:273-298)
  double dlq[12], dq1[16], G34[12];
  dlog_dq(qQ, dlq);
  { const double q2c[4] = {-s2.q[0], -s2.q[1], -s2.q[2], s2.q[3]}; dq1q2_dq1(q2c, dq1); }
  mm<3, 4, 4>(dlq, dq1, G34);          // d phi / d q_int
  double Dm[90];                        // 9 x 10
  mzero<90>(Dm);
  for (int k = 0; k < 3; ++k) { Dm[k * 10 + k] = 1.0; Dm[(6 + k) * 10 + 7 + k] = 1.0; }
  for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 4; ++c) Dm[(3 + rr) * 10 + 3 + c] = G34[rr * 4 + c];
  double cov[D * D], t90[90];
  mzero<D * D>(cov);
  for (int k = 0; k < D; ++k) cov[k * D + k] = 1.0;
  if (D >= 15) for (int k = 0; k < 6; ++k) cov[(9 + k) * D + 9 + k] = R.imu_rb[k] * total_dt;
  mm<9, 10, 10>(Dm, C, t90);
  for (int rr = 0; rr < 9; ++rr) for (int c = 0; c < 9; ++c) { double s = 0.0; for (int k = 0; k < 10; ++k) s += t90[rr * 10 + k] * Dm[c * 10 + k]; cov[rr * D + c] = s; }
  double Ci[D * D];
  invert<D>(cov, Ci);
  // bias Jacobian (parallel_algos.h:313-337)
  if (D >= 15) {
    for (int rr = 0; rr < 3; ++rr) for (int c = 0; c < 6; ++c) {
      J1[rr * D + 9 + c] = db[rr * 6 + c];
      double s = 0.0;
      for (int k = 0; k < 4; ++k) s += G34[rr * 4 + k] * db[(3 + k) * 6 + c];
      J1[(3 + rr) * D + 9 + c] = s;
      J1[(6 + rr) * D + 9 + c] = db[(7 + rr) * 6 + c];
    }
    for (int k = 0; k < 6; ++k) { J1[(9 + k) * D + 9 + k] = 1.0; J2[(9 + k) * D + 9 + k] = -1.0; }
  }
  // maha and Huber weight (cpp:1511-1526); threshold from the projection errors (Q3)
  double maha = 0.0;
  for (int rr = 0; rr < D; ++rr) for (int c = 0; c < D; ++c) maha += r[rr] * Ci[rr * D + c] * r[c];
  const int o1 = P.pose_opt[p1], o2 = P.pose_opt[p2];
  double weight = 1.0;
  const bool have_proj = P.sc->dot[7] > 0.5;
  if (have_proj) {
    const double c_huber = 1.2107 * sqrt(P.sc->median);
    const double e = sqrt(maha);
    const bool is_cond = (o1 < 0) && (o2 >= 0);
    if (e > c_huber && !is_cond && (P.use_robust & 2)) weight = c_huber / e;
  }
  double* Cout = R.imu_cov_inv + (size_t)i * D * D;
  double* J1o = R.imu_J1 + (size_t)i * D * D;
  double* J2o = R.imu_J2 + (size_t)i * D * D;
  double* ro = R.imu_res + (size_t)i * D;
  for (int k = 0; k < D * D; ++k) { Ci[k] *= weight; Cout[k] = Ci[k]; }
  // parameter masks (cpp:1698-1705)
  if (o1 >= 0) { const uint32_t fr = P.pose_free[p1]; for (int k = 0; k < D; ++k) if (!((fr >> k) & 1u)) for (int rr = 0; rr < D; ++rr) J1[rr * D + k] = 0.0; }
  if (o2 >= 0) { const uint32_t fr = P.pose_free[p2]; for (int k = 0; k < D; ++k) if (!((fr >> k) & 1u)) for (int rr = 0; rr < D; ++rr) J2[rr * D + k] = 0.0; }
  for (int k = 0; k < D * D; ++k) { J1o[k] = (o1 >= 0) ? J1[k] : 0.0; J2o[k] = (o2 >= 0) ? J2[k] : 0.0; }
  for (int k = 0; k < D; ++k) ro[k] = r[k];
  if (!have_proj) return 0.0;   // Q3: nothing is written to r_i_ / J_i without projection residuals
  // U and b_p: J^T C J, J^T C r
  double Cr[D];
  for (int rr = 0; rr < D; ++rr) { double s = 0.0; for (int c = 0; c < D; ++c) s += Ci[rr * D + c] * r[c]; Cr[rr] = s; }
  double CJ[D * D];
  for (int pass = 0; pass < 2; ++pass) {
    const int o = pass ? o2 : o1;
    if (o < 0) continue;
    const double* J = pass ? J2 : J1;
    for (int rr = 0; rr < D; ++rr) for (int c = 0; c < D; ++c) { double s = 0.0; for (int k = 0; k < D; ++k) s += Ci[rr * D + k] * J[k * D + c]; CJ[rr * D + c] = s; }
    const SBlock blk = s_block(P, o, o);
    for (int c = 0; c < D; ++c) for (int rr = 0; rr <= c; ++rr) { double s = 0.0; for (int k = 0; k < D; ++k) s += J[k * D + rr] * CJ[k * D + c]; red_add(blk.p + (size_t)rr * blk.rs + (size_t)c * blk.cs, s); }
    for (int c = 0; c < D; ++c) { double s = 0.0; for (int k = 0; k < D; ++k) s += J[k * D + c] * Cr[k]; red_add(P.b_p + (size_t)o * D + c, s); }
    if (pass == 1 && o1 >= 0 && o1 != o2) {
      // cross block (min, max): J_min^T C J_max ; CJ currently holds C J2
      const bool first1 = o1 < o2;
      if (!first1) for (int rr = 0; rr < D; ++rr) for (int c = 0; c < D; ++c) { double s = 0.0; for (int k = 0; k < D; ++k) s += Ci[rr * D + k] * J1[k * D + c]; CJ[rr * D + c] = s; }
      const double* Ja = first1 ? J1 : J2;
      const SBlock xb = first1 ? s_block(P, o1, o2) : s_block(P, o2, o1);
      for (int c = 0; c < D; ++c) for (int rr = 0; rr < D; ++rr) { double s = 0.0; for (int k = 0; k < D; ++k) s += Ja[k * D + rr] * CJ[k * D + c]; red_add(xb.p + (size_t)rr * xb.rs + (size_t)c * xb.cs, s); }
    }
  }
  return maha * weight;
}

template <int D>
BA_D double imu_sd_one(const DeviceProblem& P, const OtherResiduals& R, int i) {
  if (!(P.sc->dot[7] > 0.5)) return 0.0;
  const int o1 = P.pose_opt[R.imu_p1[i]], o2 = P.pose_opt[R.imu_p2[i]];
  const double* J1 = R.imu_J1 + (size_t)i * D * D;
  const double* J2 = R.imu_J2 + (size_t)i * D * D;
  const double* C = R.imu_cov_inv + (size_t)i * D * D;
  double v[D];
  for (int r = 0; r < D; ++r) {
    double s = 0.0;
    if (o1 >= 0) for (int k = 0; k < D; ++k) s += J1[r * D + k] * P.b_p[(size_t)o1 * D + k];
    if (o2 >= 0) for (int k = 0; k < D; ++k) s += J2[r * D + k] * P.b_p[(size_t)o2 * D + k];
    v[r] = s;
  }
  double s = 0.0;
  for (int r = 0; r < D; ++r) for (int c = 0; c < D; ++c) s += v[r] * C[r * D + c] * v[c];
  return s;
}

template <int D>
BA_D double imu_eval_one(const DeviceProblem& P, const OtherResiduals& R, int i) {
  const int p1 = (int)R.imu_p1[i], p2 = (int)R.imu_p2[i];
  ImuState s1, s2, y, yn;
  double b1[6], b2[6];
  load_state(P, p1, s1, b1); load_state(P, p2, s2, b2);
  y = s1;
  const int m0 = (int)R.imu_meas_start[i], m1 = (int)R.imu_meas_start[i + 1];
  ImuSample za, zb;
  for (int m = m0 + 1; m < m1; ++m) {
    load_sample(R.imu_meas + 7 * (size_t)(m - 1), za); load_sample(R.imu_meas + 7 * (size_t)m, zb);
    if (zb.time - za.time == 0) continue;
    integrate_imu(y, za, zb, b1, b1 + 3, R.gravity, yn);
    y = yn;
  }
  double r[D], phi[3], qQ[4];
  imu_residual<D>(y, s2, b1, b2, r, phi, qQ);
  const double* C = R.imu_cov_inv + (size_t)i * D * D;
  double s = 0.0;
  for (int rr = 0; rr < D; ++rr) for (int c = 0; c < D; ++c) s += r[rr] * C[rr * D + c] * r[c];
  double* out = R.imu_out + (size_t)i * (D + 1);
  for (int rr = 0; rr < D; ++rr) out[rr] = r[rr];
  out[D] = s;
  return s;
}

}  // namespace bagpu
