/*
 * MiniTypes.h — the few fixed-size vector / matrix / SE3 value types the host-side
 * ba::BundleAdjuster template needs at its public interface.
 *
 * The reference spells these as Eigen::Matrix<Scalar,R,C> and Sophus::SE3Group<Scalar>
 * (reference include/ba/BundleAdjuster.h:146-155).  Neither library exists in this image, so
 * the drop-in header carries value types with the same spelling of the calls user code makes
 * on them (operator[] / operator()(r,c) / data() / Zero() / Identity() / translation() /
 * unit_quaternion().coeffs() / so3() / inverse() / operator*).  A build that has Eigen and
 * Sophus defines BA_B200_HAVE_EIGEN_SOPHUS before including ba/BundleAdjuster.h and gets the
 * real types instead; only the four adaptor functions at the bottom of
 * this file touch their internals.
 *
 * Conventions: Hamilton quaternion, coefficient order x y z w, the SO3
 * constructor and the group product normalise the quaternion (Sophus semantics).
 */
#ifndef BA_B200_MINITYPES_H
#define BA_B200_MINITYPES_H

#include <cmath>
#include <cstddef>
#include <cstring>
#include <initializer_list>

namespace ba {
namespace mini {

template <typename Scalar, int R, int C = 1>
struct Matrix {
  Scalar v[R * C > 0 ? R * C : 1];   // column-major like Eigen
  Matrix() { for (int i = 0; i < R * C; ++i) v[i] = Scalar(0); }
  Matrix(std::initializer_list<Scalar> l) {
    int i = 0;
    for (Scalar x : l) { if (i < R * C) v[i++] = x; }
    for (; i < R * C; ++i) v[i] = Scalar(0);
  }
  static Matrix Zero() { return Matrix(); }
  static Matrix Identity() { Matrix m; for (int i = 0; i < (R < C ? R : C); ++i) m(i, i) = Scalar(1); return m; }
  static constexpr int rows() { return R; }
  static constexpr int cols() { return C; }
  static constexpr int size() { return R * C; }
  Scalar& operator[](int i) { return v[i]; }
  const Scalar& operator[](int i) const { return v[i]; }
  Scalar& operator()(int r, int c) { return v[c * R + r]; }
  const Scalar& operator()(int r, int c) const { return v[c * R + r]; }
  Scalar* data() { return v; }
  const Scalar* data() const { return v; }
  Matrix& setZero() { for (int i = 0; i < R * C; ++i) v[i] = Scalar(0); return *this; }
  Scalar squaredNorm() const { Scalar s = 0; for (int i = 0; i < R * C; ++i) s += v[i] * v[i]; return s; }
  Scalar norm() const { return std::sqrt(squaredNorm()); }
  Matrix operator+(const Matrix& o) const { Matrix m; for (int i = 0; i < R * C; ++i) m.v[i] = v[i] + o.v[i]; return m; }
  Matrix operator-(const Matrix& o) const { Matrix m; for (int i = 0; i < R * C; ++i) m.v[i] = v[i] - o.v[i]; return m; }
  Matrix operator*(Scalar s) const { Matrix m; for (int i = 0; i < R * C; ++i) m.v[i] = v[i] * s; return m; }
  template <int K>
  Matrix<Scalar, R, K> operator*(const Matrix<Scalar, C, K>& o) const {
    Matrix<Scalar, R, K> m;
    for (int r = 0; r < R; ++r) for (int k = 0; k < K; ++k) { Scalar s = 0; for (int c = 0; c < C; ++c) s += (*this)(r, c) * o(c, k); m(r, k) = s; }
    return m;
  }
  Matrix<Scalar, C, R> transpose() const { Matrix<Scalar, C, R> m; for (int r = 0; r < R; ++r) for (int c = 0; c < C; ++c) m(c, r) = (*this)(r, c); return m; }
  /* Gauss-Jordan with partial pivoting; the reference calls Eigen's fixed-size inverse() on 6x6
   * covariances at the API boundary (reference include/ba/BundleAdjuster.h:398,444). */
  Matrix inverse() const {
    static_assert(R == C, "inverse() of a square matrix");
    Scalar a[R][2 * R];
    for (int r = 0; r < R; ++r) for (int c = 0; c < R; ++c) { a[r][c] = (*this)(r, c); a[r][R + c] = (r == c) ? Scalar(1) : Scalar(0); }
    for (int c = 0; c < R; ++c) {
      int p = c;
      for (int r = c + 1; r < R; ++r) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
      if (p != c) for (int k = 0; k < 2 * R; ++k) { Scalar t = a[c][k]; a[c][k] = a[p][k]; a[p][k] = t; }
      const Scalar d = Scalar(1) / a[c][c];
      for (int k = 0; k < 2 * R; ++k) a[c][k] *= d;
      for (int r = 0; r < R; ++r) {
        if (r == c) continue;
        const Scalar f = a[r][c];
        if (f != Scalar(0)) for (int k = 0; k < 2 * R; ++k) a[r][k] -= f * a[c][k];
      }
    }
    Matrix m;
    for (int r = 0; r < R; ++r) for (int c = 0; c < R; ++c) m(r, c) = a[r][R + c];
    return m;
  }
};

template <typename Scalar>
struct Quaternion {
  Matrix<Scalar, 4, 1> c;   // x y z w
  Quaternion() { c[3] = Scalar(1); }
  Quaternion(Scalar w, Scalar x, Scalar y, Scalar z) { c[0] = x; c[1] = y; c[2] = z; c[3] = w; }   // Eigen's argument order
  const Matrix<Scalar, 4, 1>& coeffs() const { return c; }
  Matrix<Scalar, 4, 1>& coeffs() { return c; }
  Scalar x() const { return c[0]; } Scalar y() const { return c[1]; } Scalar z() const { return c[2]; } Scalar w() const { return c[3]; }
  void normalize() { const Scalar n = c.norm(); for (int i = 0; i < 4; ++i) c[i] /= n; }
  Quaternion conjugate() const { return Quaternion(c[3], -c[0], -c[1], -c[2]); }
  Quaternion operator*(const Quaternion& b) const {
    const Scalar ax = c[0], ay = c[1], az = c[2], aw = c[3], bx = b.c[0], by = b.c[1], bz = b.c[2], bw = b.c[3];
    return Quaternion(aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                      aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx);
  }
};

template <typename Scalar>
struct SO3Group {
  Quaternion<Scalar> q;
  SO3Group() {}
  explicit SO3Group(const Quaternion<Scalar>& qq) : q(qq) { q.normalize(); }
  const Quaternion<Scalar>& unit_quaternion() const { return q; }
  Matrix<Scalar, 3, 3> matrix() const {
    const Scalar x = q.x(), y = q.y(), z = q.z(), w = q.w();
    Matrix<Scalar, 3, 3> R;
    R(0, 0) = 1 - 2 * (y * y + z * z); R(0, 1) = 2 * (x * y - w * z); R(0, 2) = 2 * (x * z + w * y);
    R(1, 0) = 2 * (x * y + w * z); R(1, 1) = 1 - 2 * (x * x + z * z); R(1, 2) = 2 * (y * z - w * x);
    R(2, 0) = 2 * (x * z - w * y); R(2, 1) = 2 * (y * z + w * x); R(2, 2) = 1 - 2 * (x * x + y * y);
    return R;
  }
  SO3Group inverse() const { SO3Group r; r.q = q.conjugate(); return r; }
  SO3Group operator*(const SO3Group& o) const { return SO3Group(q * o.q); }
  Matrix<Scalar, 3, 1> operator*(const Matrix<Scalar, 3, 1>& p) const { return matrix() * p; }
  static SO3Group exp(const Matrix<Scalar, 3, 1>& w) {
    const Scalar th2 = w.squaredNorm(), th = std::sqrt(th2);
    Scalar k, cw;
    if (th < Scalar(1e-10)) { k = Scalar(0.5) - th2 / Scalar(48); cw = Scalar(1) - th2 / Scalar(8); }
    else { k = std::sin(th / 2) / th; cw = std::cos(th / 2); }
    return SO3Group(Quaternion<Scalar>(cw, k * w[0], k * w[1], k * w[2]));
  }
};

template <typename Scalar>
struct SE3Group {
  SO3Group<Scalar> r;
  Matrix<Scalar, 3, 1> t;
  SE3Group() {}
  SE3Group(const SO3Group<Scalar>& rr, const Matrix<Scalar, 3, 1>& tt) : r(rr), t(tt) {}
  SE3Group(const Quaternion<Scalar>& q, const Matrix<Scalar, 3, 1>& tt) : r(q), t(tt) {}
  const SO3Group<Scalar>& so3() const { return r; }
  const Quaternion<Scalar>& unit_quaternion() const { return r.q; }
  const Matrix<Scalar, 3, 1>& translation() const { return t; }
  Matrix<Scalar, 3, 1>& translation() { return t; }
  Matrix<Scalar, 3, 3> rotationMatrix() const { return r.matrix(); }
  SE3Group inverse() const { SE3Group o; o.r = r.inverse(); o.t = (o.r * t) * Scalar(-1); return o; }
  SE3Group operator*(const SE3Group& o) const { return SE3Group(r * o.r, t + r * o.t); }
  Matrix<Scalar, 3, 1> operator*(const Matrix<Scalar, 3, 1>& p) const { return r * p + t; }
};

}  // namespace mini

/* ---- adaptors: the only functions that look inside the SE3 / matrix types ------------------- */
namespace detail {
template <typename SE3>
inline void se3_to7(const SE3& T, double* o) {   /* qx qy qz qw tx ty tz */
  const auto& q = T.unit_quaternion().coeffs();
  o[0] = (double)q[0]; o[1] = (double)q[1]; o[2] = (double)q[2]; o[3] = (double)q[3];
  const auto& t = T.translation();
  o[4] = (double)t[0]; o[5] = (double)t[1]; o[6] = (double)t[2];
}
template <typename SE3, typename Scalar>
inline SE3 se3_from7(const double* p) {
  mini::Quaternion<Scalar> q((Scalar)p[3], (Scalar)p[0], (Scalar)p[1], (Scalar)p[2]);
  mini::Matrix<Scalar, 3, 1> t; t[0] = (Scalar)p[4]; t[1] = (Scalar)p[5]; t[2] = (Scalar)p[6];
  return SE3(q, t);
}
}  // namespace detail
}  // namespace ba

#endif /* BA_B200_MINITYPES_H */
