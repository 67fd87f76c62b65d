// TEST INFRASTRUCTURE — CPU oracle for the B200 BundleAdjuster path.  PARITY UNPINNED:
// the reference (arpg/ba) cannot be compiled here (Eigen3, Sophus, Calibu, TBB absent,
// reference CMakeLists.txt:49-52) and ships no golden vectors, so this restatement is
// checked only by the reference's own finite-difference strategy (reference
// include/ba/BundleAdjusterTest.h, Utils.h:68-69) and by NumPy dense re-derivations.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, link or call anything under oracle/.
//
// minimath.h: fixed-size matrices plus the Sophus / Eigen semantics the reference
// relies on (third-party, not vendored under /root/reference; version unpinned,
// pre-1.0 Sophus API `SE3Group/SO3Group`):
//   * quaternion = Hamilton product, coefficient order (x,y,z,w) (Eigen::Quaternion)
//   * SO3::exp / SO3::log with Sophus' small-angle branches (epsilon 1e-10), log uses atan
//   * SO3 group product re-normalises the quaternion
//   * SO3 * point = Eigen's _transformVector (v + w*uv + q x uv, uv = 2 q x v)
#pragma once
#include <cmath>
#include <cstring>
#include <cstdint>
#include <cassert>
#include <vector>
#include <algorithm>

namespace orc {

template <int R, int C>
struct Mat {
  double a[R * C > 0 ? R * C : 1];
  static constexpr int Rows = R, Cols = C;
  Mat() { setZero(); }
  void setZero() { for (int i = 0; i < R * C; ++i) a[i] = 0.0; }
  void setIdentity() { setZero(); for (int i = 0; i < (R < C ? R : C); ++i) (*this)(i, i) = 1.0; }
  static Mat Zero() { return Mat(); }
  static Mat Identity() { Mat m; m.setIdentity(); return m; }
  double& operator()(int i, int j) { return a[i * C + j]; }
  const double& operator()(int i, int j) const { return a[i * C + j]; }
  double& operator[](int i) { return a[i]; }             // vectors
  const double& operator[](int i) const { return a[i]; }
  Mat<C, R> transpose() const { Mat<C, R> t; for (int i = 0; i < R; ++i) for (int j = 0; j < C; ++j) t(j, i) = (*this)(i, j); return t; }
  Mat operator+(const Mat& o) const { Mat r; for (int i = 0; i < R * C; ++i) r.a[i] = a[i] + o.a[i]; return r; }
  Mat operator-(const Mat& o) const { Mat r; for (int i = 0; i < R * C; ++i) r.a[i] = a[i] - o.a[i]; return r; }
  Mat operator-() const { Mat r; for (int i = 0; i < R * C; ++i) r.a[i] = -a[i]; return r; }
  Mat operator*(double s) const { Mat r; for (int i = 0; i < R * C; ++i) r.a[i] = a[i] * s; return r; }
  Mat operator/(double s) const { Mat r; for (int i = 0; i < R * C; ++i) r.a[i] = a[i] / s; return r; }
  Mat& operator+=(const Mat& o) { for (int i = 0; i < R * C; ++i) a[i] += o.a[i]; return *this; }
  Mat& operator-=(const Mat& o) { for (int i = 0; i < R * C; ++i) a[i] -= o.a[i]; return *this; }
  Mat& operator*=(double s) { for (int i = 0; i < R * C; ++i) a[i] *= s; return *this; }
  double squaredNorm() const { double s = 0; for (int i = 0; i < R * C; ++i) s += a[i] * a[i]; return s; }
  double norm() const { return std::sqrt(squaredNorm()); }
  template <int BR, int BC> Mat<BR, BC> block(int r0, int c0) const {
    Mat<BR, BC> b; for (int i = 0; i < BR; ++i) for (int j = 0; j < BC; ++j) b(i, j) = (*this)(r0 + i, c0 + j); return b; }
  template <int BR, int BC> void setBlock(int r0, int c0, const Mat<BR, BC>& b) {
    for (int i = 0; i < BR; ++i) for (int j = 0; j < BC; ++j) (*this)(r0 + i, c0 + j) = b(i, j); }
  template <int BR, int BC> void addBlock(int r0, int c0, const Mat<BR, BC>& b) {
    for (int i = 0; i < BR; ++i) for (int j = 0; j < BC; ++j) (*this)(r0 + i, c0 + j) += b(i, j); }
  void setColZero(int c) { for (int i = 0; i < R; ++i) (*this)(i, c) = 0.0; }
};

template <int R, int K, int C>
inline Mat<R, C> operator*(const Mat<R, K>& A, const Mat<K, C>& B) {
  Mat<R, C> r;
  for (int i = 0; i < R; ++i)
    for (int j = 0; j < C; ++j) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += A(i, k) * B(k, j);
      r(i, j) = s;
    }
  return r;
}
template <int R, int C> inline Mat<R, C> operator*(double s, const Mat<R, C>& A) { return A * s; }

typedef Mat<2, 1> Vec2; typedef Mat<3, 1> Vec3; typedef Mat<4, 1> Vec4;
typedef Mat<6, 1> Vec6; typedef Mat<7, 1> Vec7; typedef Mat<3, 3> Mat3;

inline Vec3 cross(const Vec3& a, const Vec3& b) {
  Vec3 r; r[0] = a[1] * b[2] - a[2] * b[1]; r[1] = a[2] * b[0] - a[0] * b[2]; r[2] = a[0] * b[1] - a[1] * b[0]; return r; }
inline double dot(const Vec3& a, const Vec3& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline Mat3 hat(const Vec3& v) { Mat3 m; m(0, 1) = -v[2]; m(0, 2) = v[1]; m(1, 0) = v[2]; m(1, 2) = -v[0]; m(2, 0) = -v[1]; m(2, 1) = v[0]; return m; }

// Eigen::Quaternion semantics, coefficients (x,y,z,w).
struct Quat {
  double x = 0, y = 0, z = 0, w = 1;
  Quat() {}
  Quat(double w_, double x_, double y_, double z_) : x(x_), y(y_), z(z_), w(w_) {}  // Eigen ctor order
  Vec3 vec() const { Vec3 v; v[0] = x; v[1] = y; v[2] = z; return v; }
  Quat conjugate() const { return Quat(w, -x, -y, -z); }
  double squaredNorm() const { return x * x + y * y + z * z + w * w; }
  void normalize() { double n = std::sqrt(squaredNorm()); x /= n; y /= n; z /= n; w /= n; }
  Mat3 toRotationMatrix() const {
    Mat3 r;
    const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w;
    const double txx = tx * x, txy = ty * x, txz = tz * x;
    const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    r(0, 0) = 1 - (tyy + tzz); r(0, 1) = txy - twz; r(0, 2) = txz + twy;
    r(1, 0) = txy + twz; r(1, 1) = 1 - (txx + tzz); r(1, 2) = tyz - twx;
    r(2, 0) = txz - twy; r(2, 1) = tyz + twx; r(2, 2) = 1 - (txx + tyy);
    return r;
  }
  Vec3 transformVector(const Vec3& v) const {  // Eigen QuaternionBase::_transformVector
    Vec3 uv = cross(vec(), v); uv = uv + uv;
    return v + uv * w + cross(vec(), uv);
  }
};
inline Quat operator*(const Quat& a, const Quat& b) {  // Hamilton
  return Quat(a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z,
              a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
              a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
              a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x);
}

static const double kSophusEps = 1e-10;

struct SO3 {
  Quat q;
  SO3() {}
  explicit SO3(const Quat& q_) : q(q_) { q.normalize(); }   // Sophus ctor normalises
  static SO3 raw(const Quat& q_) { SO3 s; s.q = q_; return s; }  // memcpy into data() (Types.h:338-339)
  const Quat& unit_quaternion() const { return q; }
  Mat3 matrix() const { return q.toRotationMatrix(); }
  Mat3 Adj() const { return matrix(); }
  SO3 inverse() const { return raw(q.conjugate()); }
  SO3 operator*(const SO3& o) const { SO3 r; r.q = q * o.q; r.q.normalize(); return r; }
  Vec3 operator*(const Vec3& p) const { return q.transformVector(p); }
  static SO3 exp(const Vec3& omega) {
    const double theta = omega.norm();
    const double half_theta = 0.5 * theta;
    double imag_factor;
    const double real_factor = std::cos(half_theta);
    if (theta < kSophusEps) {
      const double theta_sq = theta * theta;
      const double theta_po4 = theta_sq * theta_sq;
      imag_factor = 0.5 - 0.0208333 * theta_sq + 0.000260417 * theta_po4;
    } else {
      imag_factor = std::sin(half_theta) / theta;
    }
    return SO3(Quat(real_factor, imag_factor * omega[0], imag_factor * omega[1], imag_factor * omega[2]));
  }
  Vec3 log() const {
    const double squared_n = q.x * q.x + q.y * q.y + q.z * q.z;
    const double n = std::sqrt(squared_n);
    const double w = q.w;
    double two_atan_nbyw_by_n;
    if (n < kSophusEps) {
      const double squared_w = w * w;
      two_atan_nbyw_by_n = 2.0 / w - 2.0 * squared_n / (w * squared_w);
    } else {
      if (std::fabs(w) < kSophusEps) {
        two_atan_nbyw_by_n = (w > 0 ? M_PI : -M_PI) / n;
      } else {
        two_atan_nbyw_by_n = 2.0 * std::atan(n / w) / n;
      }
    }
    return q.vec() * two_atan_nbyw_by_n;
  }
  static Mat3 generator(int i) { Vec3 e; e[i] = 1.0; return hat(e); }
};

struct SE3 {
  SO3 r;
  Vec3 t;
  SE3() {}
  SE3(const SO3& r_, const Vec3& t_) : r(r_), t(t_) {}
  const SO3& so3() const { return r; }
  SO3& so3() { return r; }
  const Quat& unit_quaternion() const { return r.q; }
  const Vec3& translation() const { return t; }
  Vec3& translation() { return t; }
  Mat3 rotationMatrix() const { return r.matrix(); }
  SE3 inverse() const { SO3 ri = r.inverse(); return SE3(ri, ri * (t * -1.0)); }
  SE3 operator*(const SE3& o) const { return SE3(r * o.r, t + r * o.t); }
  Mat<4, 4> matrix() const {
    Mat<4, 4> m; m.setBlock(0, 0, r.matrix()); m.setBlock(0, 3, t); m(3, 3) = 1.0; return m; }
  static SE3 from7(const double* p) {  // qx qy qz qw tx ty tz
    SE3 T; T.r = SO3(Quat(p[3], p[0], p[1], p[2])); T.t[0] = p[4]; T.t[1] = p[5]; T.t[2] = p[6]; return T; }
  void to7(double* p) const { p[0] = r.q.x; p[1] = r.q.y; p[2] = r.q.z; p[3] = r.q.w; p[4] = t[0]; p[5] = t[1]; p[6] = t[2]; }
};

// ---- dynamic helpers -------------------------------------------------------------
typedef std::vector<double> VecX;
inline double squaredNorm(const VecX& v) { double s = 0; for (double x : v) s += x * x; return s; }

// General small inverse (Gauss-Jordan, partial pivoting) standing in for Eigen's
// fixed-size inverse() on 6x6 / 9x9 / 15x15 (reference BundleAdjuster.h:398,444, parallel_algos.h:307).
template <int N>
inline Mat<N, N> inverse(const Mat<N, N>& A) {
  Mat<N, N> M = A, I = Mat<N, N>::Identity();
  for (int c = 0; c < N; ++c) {
    int p = c; double best = std::fabs(M(c, c));
    for (int r = c + 1; r < N; ++r) if (std::fabs(M(r, c)) > best) { best = std::fabs(M(r, c)); p = r; }
    if (p != c) for (int j = 0; j < N; ++j) { std::swap(M(c, j), M(p, j)); std::swap(I(c, j), I(p, j)); }
    const double d = 1.0 / M(c, c);
    for (int j = 0; j < N; ++j) { M(c, j) *= d; I(c, j) *= d; }
    for (int r = 0; r < N; ++r) if (r != c) {
      const double f = M(r, c);
      if (f != 0.0) for (int j = 0; j < N; ++j) { M(r, j) -= f * M(c, j); I(r, j) -= f * I(c, j); }
    }
  }
  return I;
}
// Eigen closed forms for 1x1 and 3x3 (cofactors), reference src/BundleAdjuster.cpp:442.
inline Mat<1, 1> inverse(const Mat<1, 1>& A) { Mat<1, 1> r; r(0, 0) = 1.0 / A(0, 0); return r; }
inline Mat<3, 3> inverse(const Mat<3, 3>& m) {
  Mat<3, 3> c;
  c(0, 0) = m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1);
  c(0, 1) = m(0, 2) * m(2, 1) - m(0, 1) * m(2, 2);
  c(0, 2) = m(0, 1) * m(1, 2) - m(0, 2) * m(1, 1);
  c(1, 0) = m(1, 2) * m(2, 0) - m(1, 0) * m(2, 2);
  c(1, 1) = m(0, 0) * m(2, 2) - m(0, 2) * m(2, 0);
  c(1, 2) = m(0, 2) * m(1, 0) - m(0, 0) * m(1, 2);
  c(2, 0) = m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0);
  c(2, 1) = m(0, 1) * m(2, 0) - m(0, 0) * m(2, 1);
  c(2, 2) = m(0, 0) * m(1, 1) - m(0, 1) * m(1, 0);
  const double det = m(0, 0) * c(0, 0) + m(0, 1) * c(1, 0) + m(0, 2) * c(2, 0);
  return c * (1.0 / det);
}

// Principal square root of a symmetric positive (semi)definite matrix by cyclic Jacobi
// eigen-decomposition; stands in for Eigen's MatrixFunctions sqrt() (reference
// BundleAdjuster.h:399,445, src/BundleAdjuster.cpp:1470,1527) — equal for SPD input.
template <int N>
inline Mat<N, N> sqrtm_spd(const Mat<N, N>& Ain) {
  Mat<N, N> A = (Ain + Ain.transpose()) * 0.5, V = Mat<N, N>::Identity();
  for (int sweep = 0; sweep < 100; ++sweep) {
    double off = 0, diag = 0;
    for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) { if (i != j) off += A(i, j) * A(i, j); else diag += A(i, j) * A(i, j); }
    if (off <= 1e-40 * diag || off == 0.0) break;
    for (int p = 0; p < N - 1; ++p)
      for (int q = p + 1; q < N; ++q) {
        if (A(p, q) == 0.0) continue;
        const double theta = (A(q, q) - A(p, p)) / (2.0 * A(p, q));
        const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < N; ++k) { const double akp = A(k, p), akq = A(k, q); A(k, p) = c * akp - s * akq; A(k, q) = s * akp + c * akq; }
        for (int k = 0; k < N; ++k) { const double apk = A(p, k), aqk = A(q, k); A(p, k) = c * apk - s * aqk; A(q, k) = s * apk + c * aqk; }
        for (int k = 0; k < N; ++k) { const double vkp = V(k, p), vkq = V(k, q); V(k, p) = c * vkp - s * vkq; V(k, q) = s * vkp + c * vkq; }
      }
  }
  Mat<N, N> R;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) {
      double s = 0;
      for (int k = 0; k < N; ++k) s += V(i, k) * std::sqrt(A(k, k) > 0 ? A(k, k) : 0.0) * V(j, k);
      R(i, j) = s;
    }
  return R;
}

}  // namespace orc
