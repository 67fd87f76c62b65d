// TEST INFRASTRUCTURE — CPU oracle (see minimath.h header; PARITY UNPINNED).
// Block-CSC container and products restating reference include/ba/SparseBlockMatrix.h
// (insert / forceTranspose semantics, SparseBlockMatrix.h:175-281,316-451) and
// include/ba/SparseBlockMatrixOps.h.  Output patterns follow the reference's
// BlockAmbiVector accumulator in sparse mode (BlockAmbiVector.h:174-241): per output
// column the inner indices come out sorted ascending and explicit zero blocks are kept.
#pragma once
#include "minimath.h"
#include <vector>

namespace orc {

template <int R, int C>
struct BlockMat {
  typedef Mat<R, C> Block;
  struct Entry { int index; Block value; };
  int nrows = 0, ncols = 0;
  std::vector<std::vector<Entry>> col;  // per outer (column): entries with ascending inner index

  BlockMat() {}
  BlockMat(int r, int c) { resize(r, c); }
  void resize(int r, int c) { nrows = r; ncols = c; col.assign(c, std::vector<Entry>()); }
  void setZero() { for (auto& v : col) v.clear(); }   // removes all blocks (SparseBlockMatrix.h setZero)
  int rows() const { return nrows; }
  int cols() const { return ncols; }
  size_t nonZeros() const { size_t n = 0; for (auto& v : col) n += v.size(); return n; }
  void reserve(int j, size_t n) { col[j].reserve(n); }
  // SparseBlockMatrix::insert(row, col): new zero block, kept sorted by inner index.
  Block& insert(int i, int j) {
    std::vector<Entry>& v = col[j];
    size_t p = v.size();
    while (p > 0 && v[p - 1].index > i) --p;
    assert(p == 0 || v[p - 1].index != i);
    Entry e; e.index = i;
    v.insert(v.begin() + p, e);
    return v[p].value;
  }
  void insertBack(int i, int j, const Block& b) { Entry e; e.index = i; e.value = b; col[j].push_back(e); }
  const Block* find(int i, int j) const {
    const std::vector<Entry>& v = col[j];
    size_t lo = 0, hi = v.size();
    while (lo < hi) { size_t mid = (lo + hi) / 2; if (v[mid].index < i) lo = mid + 1; else hi = mid; }
    return (lo < v.size() && v[lo].index == i) ? &v[lo].value : nullptr;
  }
  Block coeff(int i, int j) const { const Block* b = find(i, j); return b ? *b : Block(); }
  // SparseBlockMatrix.h:234-281: dest(j,i) = src(i,j)^T
  static void forceTranspose(const BlockMat<C, R>& src, BlockMat<R, C>& dest) {
    dest.resize(src.cols(), src.rows());
    for (int j = 0; j < src.cols(); ++j)
      for (const auto& e : src.col[j]) dest.insertBack(j, e.index, e.value.transpose());
  }
  void pattern(std::vector<uint32_t>& colptr, std::vector<uint32_t>& rowidx) const {
    colptr.assign(ncols + 1, 0); rowidx.clear();
    for (int j = 0; j < ncols; ++j) { for (const auto& e : col[j]) rowidx.push_back(e.index); colptr[j + 1] = (uint32_t)rowidx.size(); }
  }
};

// SparseBlockMatrixOps.h:6-38
template <int R, int C>
inline void SparseBlockVectorProductDenseResult(const BlockMat<R, C>& lhs, const VecX& rhs, VecX& res,
                                                int rhs_stride = -1, int res_stride = -1) {
  if (res_stride == -1) res_stride = R;
  if (rhs_stride == -1) rhs_stride = C;
  std::fill(res.begin(), res.end(), 0.0);
  for (int ii = 0; ii < lhs.cols(); ++ii)
    for (const auto& e : lhs.col[ii])
      for (int r = 0; r < R; ++r) {
        double s = 0;
        for (int c = 0; c < C; ++c) s += e.value(r, c) * rhs[rhs_stride * ii + c];
        res[res_stride * e.index + r] += s;
      }
}

// SparseBlockMatrixOps.h:159-180: res = lhs * blockdiag(rhs)
template <int R, int C>
inline void SparseBlockDiagonalRhsProduct(const BlockMat<R, C>& lhs, const BlockMat<C, C>& rhs, BlockMat<R, C>& res) {
  res.resize(lhs.rows(), rhs.cols());
  for (int jj = 0; jj < rhs.cols(); ++jj) {
    const Mat<C, C> d = rhs.coeff(jj, jj);
    for (const auto& e : lhs.col[jj]) res.insertBack(e.index, jj, e.value * d);
  }
}

// SparseBlockMatrixOps.h:182-254
template <int R, int K, int C>
inline void SparseBlockProduct(const BlockMat<R, K>& lhs, const BlockMat<K, C>& rhs, BlockMat<R, C>& res,
                               const bool upper_triangular = false) {
  const int rows = lhs.rows(), cols = rhs.cols();
  res.resize(rows, cols);
  std::vector<Mat<R, C>> acc(rows);
  std::vector<char> used(rows, 0);
  std::vector<int> touched;
  int count = 0;
  for (int jj = 0; jj < cols; ++jj) {
    touched.clear();
    for (const auto& re : rhs.col[jj]) {
      for (const auto& le : lhs.col[re.index]) {
        if (upper_triangular && le.index > count) break;
        if (!used[le.index]) { used[le.index] = 1; acc[le.index].setZero(); touched.push_back(le.index); }
        acc[le.index] += le.value * re.value;
      }
    }
    count++;
    std::sort(touched.begin(), touched.end());
    for (int i : touched) { res.insertBack(i, jj, acc[i]); used[i] = 0; }
  }
}

// SparseBlockMatrixOps.h:257-316: res = lhs + rhs_coef * rhs, rhs block added at (0,0) of the lhs block type
template <int R, int C, int R2, int C2>
inline void SparseBlockAdd(const BlockMat<R, C>& lhs, const BlockMat<R2, C2>& rhs, BlockMat<R, C>& res,
                           const int rhs_coef = 1) {
  const int rows = lhs.rows(), cols = lhs.cols();
  res.resize(rows, cols);
  std::vector<Mat<R, C>> acc(rows);
  std::vector<char> used(rows, 0);
  std::vector<int> touched;
  for (int jj = 0; jj < cols; ++jj) {
    touched.clear();
    for (const auto& re : rhs.col[jj]) {
      if (!used[re.index]) { used[re.index] = 1; acc[re.index].setZero(); touched.push_back(re.index); }
      acc[re.index].template addBlock<R2, C2>(0, 0, re.value * (double)rhs_coef);
    }
    for (const auto& le : lhs.col[jj]) {
      if (!used[le.index]) { used[le.index] = 1; acc[le.index].setZero(); touched.push_back(le.index); }
      acc[le.index] += le.value;
    }
    std::sort(touched.begin(), touched.end());
    for (int i : touched) { res.insertBack(i, jj, acc[i]); used[i] = 0; }
  }
}

// Upper-triangular symmetric matrix in column-profile ("skyline") storage standing in for the
// reference's dense s_ (BundleAdjuster.h:718; dense (P*d)^2 does not fit for configs 4/5).
// Entries outside the profile are structural zeros of the reference's dense S; writes
// below the diagonal (use_triangular_matrices=false) are dropped because the LDLT only
// reads the upper triangle (src/BundleAdjuster.cpp:753,793).
struct Skyline {
  int n = 0;
  std::vector<int> first;        // first stored row of each column
  std::vector<size_t> start;     // offset of column j (row first[j]) in val
  std::vector<double> val;
  void init(int n_, const std::vector<int>& first_) {
    n = n_; first = first_; start.assign(n + 1, 0);
    for (int j = 0; j < n; ++j) start[j + 1] = start[j] + (size_t)(j - first[j] + 1);
    val.assign(start[n], 0.0);
  }
  inline double& at(int i, int j) { return val[start[j] + (size_t)(i - first[j])]; }
  inline double get(int i, int j) const { if (i > j) std::swap(i, j); return i < first[j] ? 0.0 : val[start[j] + (size_t)(i - first[j])]; }
  template <int R, int C> void addBlock(int r0, int c0, const Mat<R, C>& b, double coef) {
    for (int j = 0; j < C; ++j) for (int i = 0; i < R; ++i) { const int gi = r0 + i, gj = c0 + j; if (gi <= gj && gi >= first[gj]) at(gi, gj) += coef * b(i, j); }
  }
};

// Active-column LDL^T (A = U^T D U, U unit upper) and solve; stands in for
// Eigen::SimplicialLDLT<SparseMatrix, Upper> (src/BundleAdjuster.cpp:753-761).
// Returns false on an exactly-zero pivot (Eigen: NumericalIssue -> FactorizationError).
inline bool skyline_ldlt_solve(Skyline& A, const VecX& rhs, VecX& x) {
  const int n = A.n;
  std::vector<double> d(n, 0.0);
  for (int j = 0; j < n; ++j) {
    const int fj = A.first[j];
    double* cj = &A.val[A.start[j]];               // cj[i - fj] = a_ij -> g_ij -> u_ij
    for (int i = fj + 1; i < j; ++i) {
      const int fi = A.first[i];
      const int k0 = fi > fj ? fi : fj;
      const double* ci = &A.val[A.start[i]];
      double s = 0;
      for (int k = k0; k < i; ++k) s += ci[k - fi] * cj[k - fj];
      cj[i - fj] -= s;
    }
    double dj = cj[j - fj];
    for (int i = fj; i < j; ++i) {
      const double g = cj[i - fj];
      const double u = g / d[i];
      cj[i - fj] = u;
      dj -= u * g;
    }
    if (dj == 0.0 || dj != dj) return false;
    d[j] = dj; cj[j - fj] = dj;
  }
  x = rhs;
  for (int j = 0; j < n; ++j) { const int fj = A.first[j]; const double* cj = &A.val[A.start[j]]; double s = 0; for (int i = fj; i < j; ++i) s += cj[i - fj] * x[i]; x[j] -= s; }
  for (int j = 0; j < n; ++j) x[j] /= d[j];
  for (int j = n - 1; j >= 0; --j) { const int fj = A.first[j]; const double* cj = &A.val[A.start[j]]; const double xj = x[j]; for (int i = fj; i < j; ++i) x[i] -= cj[i - fj] * xj; }
  return true;
}

}  // namespace orc
