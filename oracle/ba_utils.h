// TEST INFRASTRUCTURE — CPU oracle (see minimath.h header; PARITY UNPINNED).
// Restatement of the SE3 / quaternion Jacobian helpers of reference include/ba/Utils.h.
// Each function cites the lines it follows.  Operand order is kept literal on purpose:
// the GPU path uses algebraically collapsed closed forms, so the two differ structurally.
#pragma once
#include "minimath.h"

namespace orc {

// Utils.h:71-82
inline Vec4 MultHomogeneous(const SE3& lhs, const Vec4& rhs) {
  Vec4 out;
  Vec3 h; h[0] = rhs[0]; h[1] = rhs[1]; h[2] = rhs[2];
  Vec3 o = lhs.so3() * h + lhs.translation() * rhs[3];
  out[0] = o[0]; out[1] = o[1]; out[2] = o[2]; out[3] = rhs[3];
  return out;
}

// Utils.h:85-100
inline double powi(const double x, const int y) {
  if (y == 0) return 1.0;
  if (y < 0) return 1.0 / powi(x, -y);
  double ret = x;
  for (int ii = 1; ii < y; ii++) ret *= x;
  return ret;
}

// Utils.h:136-185
inline Mat<3, 4> dlog_dq(const Quat& q) {
  const double x = q.x, y = q.y, z = q.z, w = q.w;
  const double vec_squarednorm = powi(x, 2) + powi(y, 2) + powi(z, 2);
  const double vec_norm = std::sqrt(vec_squarednorm);
  Mat<3, 4> m;
  if (vec_norm < 1e-9) {
    const double s1 = 2 * vec_squarednorm;
    const double s2 = 1.0 / powi(w, 3);
    const double s3 = (3 * s1) / powi(w, 4) - 2 / powi(w, 2);
    const double s4 = 2 / w;
    const double v[12] = {
        -4 * s2 * powi(x, 2) + s4 - s1 * s2, -4 * x * y * s2, -4 * x * z * s2, x * s3,
        -4 * x * y * s2, -4 * s2 * powi(y, 2) + s4 - s1 * s2, -4 * y * z * s2, y * s3,
        -4 * x * z * s2, -4 * y * z * s2, -4 * s2 * powi(z, 2) + s4 - s1 * s2, z * s3};
    for (int i = 0; i < 12; ++i) m.a[i] = v[i];
  } else {
    const double s1 = vec_squarednorm;
    const double s2 = 1 / (s1 / powi(w, 2) + 1);
    const double s3 = std::atan(std::sqrt(s1) / w);
    const double s4 = 1 / std::pow(s1, (3.0 / 2.0));
    const double s5 = 1 / s1;
    const double s6 = 1 / w;
    const double s7 = (2 * s3) / std::sqrt(s1);
    const double s8 = 2 * y * z * s2 * s5 * s6 - 2 * y * z * s3 * s4;
    const double s9 = 2 * x * z * s2 * s5 * s6 - 2 * x * z * s3 * s4;
    const double s10 = 2 * x * y * s2 * s5 * s6 - 2 * x * y * s3 * s4;
    const double v[12] = {
        s7 - 2 * powi(x, 2) * s3 * s4 + 2 * powi(x, 2) * s2 * s5 * s6, s10, s9, -(2 * x * s2) / powi(w, 2),
        s10, s7 - 2 * powi(y, 2) * s3 * s4 + 2 * powi(y, 2) * s2 * s5 * s6, s8, -(2 * y * s2) / powi(w, 2),
        s9, s8, s7 - 2 * powi(z, 2) * s3 * s4 + 2 * powi(z, 2) * s2 * s5 * s6, -(2 * z * s2) / powi(w, 2)};
    for (int i = 0; i < 12; ++i) m.a[i] = v[i];
  }
  return m;
}

// Utils.h:251-266
inline Mat<4, 3> dq_exp_dw(const Vec3& w) {
  const double t = w.norm();
  const double s1 = t / 20 - 1;
  const double s2 = powi(t, 2) / 48 - 0.5;
  const double s3 = (s1 * w[1] * w[2]) / 24;
  const double s4 = (s1 * w[0] * w[2]) / 24;
  const double s5 = (s1 * w[0] * w[1]) / 24;
  const double s6 = powi(t, 2);
  Mat<4, 3> m;
  const double v[12] = {(s1 * powi(w[0], 2)) / 24 - s6 / 48 + 0.5, s5, s4,
                        s5, (s1 * powi(w[1], 2)) / 24 - s6 / 48 + 0.5, s3,
                        s4, s3, (s1 * powi(w[2], 2)) / 24 - s6 / 48 + 0.5,
                        (s2 * w[0]) / 2, (s2 * w[1]) / 2, (s2 * w[2]) / 2};
  for (int i = 0; i < 12; ++i) m.a[i] = v[i];
  return m;
}

// Utils.h:269-273
inline Mat<4, 4> dqinv_dq() { Mat<4, 4> m; m(0, 0) = -1; m(1, 1) = -1; m(2, 2) = -1; m(3, 3) = 1; return m; }

// Utils.h:276-282
inline Mat<4, 4> dq1q2_dq2(const Quat& q1) {
  Mat<4, 4> m;
  const double v[16] = {q1.w, -q1.z, q1.y, q1.x, q1.z, q1.w, -q1.x, q1.y,
                        -q1.y, q1.x, q1.w, q1.z, -q1.x, -q1.y, -q1.z, q1.w};
  for (int i = 0; i < 16; ++i) m.a[i] = v[i];
  return m;
}

// Utils.h:285-291
inline Mat<4, 4> dq1q2_dq1(const Quat& q2) {
  Mat<4, 4> m;
  const double v[16] = {q2.w, q2.z, -q2.y, q2.x, -q2.z, q2.w, q2.x, q2.y,
                        q2.y, -q2.x, q2.w, q2.z, -q2.x, -q2.y, -q2.z, q2.w};
  for (int i = 0; i < 16; ++i) m.a[i] = v[i];
  return m;
}

// Utils.h:294-312
inline Mat<3, 4> dqx_dq(const Quat& q, const Vec3& vec) {
  const double x = vec[0], y = vec[1], z = vec[2];
  const double s1 = 2 * q.x * y;
  const double s2 = 2 * q.y * y;
  const double s3 = 2 * q.x * x;
  const double s4 = 2 * q.z * x;
  const double s5 = 2 * q.y * z;
  const double s6 = 2 * q.z * z;
  Mat<3, 4> m;
  const double v[12] = {
      s2 + s6, s1 - 4 * q.y * x + 2 * q.w * z, 2 * q.x * z - 2 * q.w * y - 4 * q.z * x, s5 - 2 * q.z * y,
      2 * q.y * x - 4 * q.x * y - 2 * q.w * z, s3 + s6, s5 + 2 * q.w * x - 4 * q.z * y, s4 - 2 * q.x * z,
      s4 + 2 * q.w * y - 4 * q.x * z, 2 * q.z * y - 2 * q.w * x - 4 * q.y * z, s2 + s3, s1 - 2 * q.y * x};
  for (int i = 0; i < 12; ++i) m.a[i] = v[i];
  return m;
}

// Utils.h:353-360
inline Vec6 log_decoupled(const SE3& a, const SE3& b) {
  Vec6 res;
  Vec3 dt = a.translation() - b.translation();
  Vec3 w = (a.so3() * b.so3().inverse()).log();
  for (int i = 0; i < 3; ++i) { res[i] = dt[i]; res[3 + i] = w[i]; }
  return res;
}

// Utils.h:363-369
inline SE3 exp_decoupled(const SE3& a, const Vec6& x) {
  Vec3 xt, xw;
  for (int i = 0; i < 3; ++i) { xt[i] = x[i]; xw[i] = x[3 + i]; }
  return SE3(a.so3() * SO3::exp(xw), a.translation() + xt);
}

// Utils.h:373-384
inline Mat<6, 6> dlog_decoupled_dx(const SE3& a, const SE3& b) {
  Mat<6, 6> d = Mat<6, 6>::Identity();
  Mat<3, 3> blk = dlog_dq((a * b.inverse()).unit_quaternion()) * dq1q2_dq2(a.unit_quaternion()) *
                  dq1q2_dq1(b.inverse().unit_quaternion()) * dq_exp_dw(Vec3());
  d.setBlock(3, 3, blk);
  return d;
}

// Utils.h:387-397
inline Mat<6, 7> dLog_decoupled_dt1(const SE3& t1, const SE3& t2) {
  Mat<6, 7> d;
  d.setBlock(0, 0, Mat3::Identity());
  Mat<3, 4> blk = dlog_dq((t1 * t2.inverse()).unit_quaternion()) * dq1q2_dq1(t2.inverse().unit_quaternion());
  d.setBlock(3, 3, blk);
  return d;
}

// Utils.h:400-447
inline Mat<6, 7> dlog_decoupled_dt2(const SE3& t1, const SE3& t2) {
  Mat<6, 7> d;
  Quat qlog = (t1.so3() * t2.so3().inverse()).unit_quaternion();
  d.setBlock(0, 0, -Mat3::Identity());
  Mat<3, 4> blk = dlog_dq(qlog) * dq1q2_dq2(t1.unit_quaternion()) * dqinv_dq();
  d.setBlock(3, 3, blk);
  return d;
}

// Utils.h:450-489
inline Mat<7, 6> dexp_decoupled_dx(const SE3& t) {
  Mat<7, 6> d;
  d.setBlock(0, 0, Mat3::Identity());
  Mat<4, 3> blk = dq1q2_dq2(t.unit_quaternion()) * dq_exp_dw(Vec3());
  d.setBlock(3, 3, blk);
  return d;
}

// Utils.h:492-536
inline Mat<7, 6> dinv_exp_decoupled_dx(const SE3& t) {
  const Mat<4, 3> dq_exp = dq_exp_dw(Vec3());
  const Quat qt_inv = t.so3().inverse().unit_quaternion();
  const Mat<4, 4> dq1q2_dq_exp = dq1q2_dq1(qt_inv);
  Mat<7, 6> d;
  d.setBlock(0, 0, -t.so3().inverse().matrix());
  Mat<3, 3> b03 = dqx_dq(qt_inv, t.translation()) * dq1q2_dq_exp * dq_exp;
  d.setBlock(0, 3, b03);
  Mat<4, 3> b33 = dq1q2_dq_exp * (-dq_exp);
  d.setBlock(3, 3, b33);
  return d;
}

// Utils.h:539-583
inline Mat<4, 7> dt_x_dt(const SE3& t, const Vec4& x) {
  Mat<4, 7> d;
  d.setBlock(0, 0, Mat3::Identity() * x[3]);
  Vec3 h; h[0] = x[0]; h[1] = x[1]; h[2] = x[2];
  d.setBlock(0, 3, dqx_dq(t.unit_quaternion(), h));
  return d;
}

// Utils.h:586-639
inline Mat<7, 7> dt1_t2_dt1(const SE3& t1, const SE3& t2) {
  Mat<7, 7> d;
  d.setBlock(0, 0, Mat3::Identity());
  d.setBlock(0, 3, dqx_dq(t1.unit_quaternion(), t2.translation()));
  d.setBlock(3, 3, dq1q2_dq1(t2.unit_quaternion()));
  return d;
}

// Utils.h:642-694
inline Mat<7, 7> dt1_t2_dt2(const SE3& t1) {
  Mat<7, 7> d;
  d.setBlock(0, 0, t1.rotationMatrix());
  d.setBlock(3, 3, dq1q2_dq2(t1.unit_quaternion()));
  return d;
}

// ---- Calibu 0.1 linear camera (third-party, not under /root/reference; recollected
// published behaviour): params (fx, fy, u0, v0);
// Project(P) = (fx X/Z + u0, fy Y/Z + v0); Transfer3d(T, ray, rho) = Project(R ray + rho t);
// dTransfer3d_dray(T, ray, rho) = [dProject(P) R , dProject(P) t].
struct LinearCamera {
  double fx = 1, fy = 1, u0 = 0, v0 = 0;
  SE3 t_vs;  // CameraInterface::Pose()
  Vec2 Project(const Vec3& P) const { Vec2 p; p[0] = fx * P[0] / P[2] + u0; p[1] = fy * P[1] / P[2] + v0; return p; }
  Mat<2, 3> dProject_dray(const Vec3& P) const {
    Mat<2, 3> J;
    J(0, 0) = fx / P[2]; J(0, 2) = -fx * P[0] / (P[2] * P[2]);
    J(1, 1) = fy / P[2]; J(1, 2) = -fy * P[1] / (P[2] * P[2]);
    return J;
  }
  Vec2 Transfer3d(const SE3& t_ba, const Vec3& ray, const double rho) const {
    return Project(t_ba.rotationMatrix() * ray + t_ba.translation() * rho);
  }
  Mat<2, 4> dTransfer3d_dray(const SE3& t_ba, const Vec3& ray, const double rho) const {
    const Mat3 R = t_ba.rotationMatrix();
    const Vec3 P = R * ray + t_ba.translation() * rho;
    const Mat<2, 3> dp = dProject_dray(P);
    Mat<2, 4> J;
    J.setBlock(0, 0, dp * R);
    J.setBlock(0, 3, dp * t_ba.translation());
    return J;
  }
};

}  // namespace orc
