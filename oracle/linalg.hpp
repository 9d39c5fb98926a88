// oracle/linalg.hpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Eigen-free dense kernels used by the CPU restatement of aligator's gar solver.
// Eigen 3.4.0 (pinned in /root/reference/pixi.lock, `eigen-3.4.0-h00ab1b0_0`) is a
// third-party dependency that is NOT vendored under /root/reference and not installed
// here; its routines are restated from their published (LAPACK-equivalent) algorithms:
//   PartialPivLU  -> dgetf2-style right-looking LU with row partial pivoting
//   LLT           -> dpotf2-style lower Cholesky
//   triangular solveInPlace / GEMM / GEMV -> textbook loops
// The Bunch-Kaufman factorisation IS in the reference tree and is restated from
//   /root/reference/include/aligator/core/bunchkaufman.hpp:20-167 (unblocked factor),
//   :389-415 (sub-diagonal extraction + row swaps), :447-516 (solve in place).
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstring>
#include <vector>

namespace gar_oracle {

/// Column-major dense matrix (a vector is an n x 1 matrix).
struct Mat {
  int r = 0, c = 0;
  std::vector<double> a;
  Mat() = default;
  Mat(int r_, int c_ = 1) : r(r_), c(c_), a((size_t)r_ * c_, 0.0) {}
  double &operator()(int i, int j) { return a[(size_t)j * r + i]; }
  double operator()(int i, int j) const { return a[(size_t)j * r + i]; }
  double &operator[](int i) { return a[i]; }
  double operator[](int i) const { return a[i]; }
  double *data() { return a.data(); }
  const double *data() const { return a.data(); }
  double *col(int j) { return a.data() + (size_t)j * r; }
  const double *col(int j) const { return a.data() + (size_t)j * r; }
  int rows() const { return r; }
  int cols() const { return c; }
  void setZero() { std::fill(a.begin(), a.end(), 0.0); }
  void resize(int r_, int c_ = 1) { r = r_; c = c_; a.assign((size_t)r_ * c_, 0.0); }
  void load(const double *src) { std::memcpy(a.data(), src, sizeof(double) * a.size()); }
  void store(double *dst) const { std::memcpy(dst, a.data(), sizeof(double) * a.size()); }
};

// ---------------------------------------------------------------------------------
// GEMM family.  C (m x n) = beta*C + alpha * op(A) * op(B); column-major.
// Loop orders are chosen so the innermost loop is unit-stride (auto-vectorised).
// ---------------------------------------------------------------------------------
// Register-blocked micro-kernel: an MR x NR tile of C stays in registers across the whole k loop
// (the textbook axpy form re-loaded and re-stored a column of C for every multiply-add and ran at
// ~6 GFLOP/s per core; Eigen's kernels, which the reference uses, are register blocked as well).
// A: m x k column-major (lda), Bt(l, j) = element (l, j) of op(B) read through (ldb_l, ldb_j) strides.
template <int MR, int NR>
static inline void gemm_tile(int kk, double alpha, const double *A, int lda, const double *B, long sl, long sj,
                             double beta, double *C, int ldc, int mr, int nr) {
  double acc[NR][MR];
  for (int j = 0; j < NR; j++)
    for (int i = 0; i < MR; i++) acc[j][i] = 0.0;
  if (mr == MR && nr == NR) {
    for (int l = 0; l < kk; l++) {
      const double *al = A + (size_t)l * lda;
      for (int j = 0; j < NR; j++) {
        const double b = B[l * sl + j * sj];
#pragma omp simd
        for (int i = 0; i < MR; i++) acc[j][i] += al[i] * b;
      }
    }
  } else {
    for (int l = 0; l < kk; l++) {
      const double *al = A + (size_t)l * lda;
      for (int j = 0; j < nr; j++) {
        const double b = B[l * sl + j * sj];
        for (int i = 0; i < mr; i++) acc[j][i] += al[i] * b;
      }
    }
  }
  for (int j = 0; j < nr; j++) {
    double *cj = C + (size_t)j * ldc;
    if (beta == 0.0) for (int i = 0; i < mr; i++) cj[i] = alpha * acc[j][i];
    else for (int i = 0; i < mr; i++) cj[i] = beta * cj[i] + alpha * acc[j][i];
  }
}
// C (m x n) = beta C + alpha A (m x k, column-major) * op(B), op(B)(l, j) = B[l * sl + j * sj]
static inline void gemm_core(int m, int n, int kk, double alpha, const double *A, int lda, const double *B, long sl,
                             long sj, double beta, double *C, int ldc) {
  constexpr int MR = 16, NR = 4;
  for (int j0 = 0; j0 < n; j0 += NR)
    for (int i0 = 0; i0 < m; i0 += MR)
      gemm_tile<MR, NR>(kk, alpha, A + i0, lda, B + j0 * sj, sl, sj, beta, C + i0 + (size_t)j0 * ldc, ldc,
                        m - i0 < MR ? m - i0 : MR, n - j0 < NR ? n - j0 : NR);
}
// scratch for the transposed-A forms (no allocation after the first call on a thread)
static inline double *tls_scratch(size_t n) {
  static thread_local std::vector<double> buf;
  if (buf.size() < n) buf.resize(n);
  return buf.data();
}
inline void gemm_nn(double alpha, const Mat &A, const Mat &B, double beta, Mat &C) {
  const int m = A.r, k = A.c, n = B.c;
  assert(B.r == k && C.r == m && C.c == n);
  if (m == 0 || n == 0) return;
  gemm_core(m, n, k, alpha, A.a.data(), m, B.a.data(), 1, k, beta, C.a.data(), m);
}
/// C = beta*C + alpha * A^T * B
inline void gemm_tn(double alpha, const Mat &A, const Mat &B, double beta, Mat &C) {
  const int m = A.c, k = A.r, n = B.c;
  assert(B.r == k && C.r == m && C.c == n);
  if (m == 0 || n == 0) return;
  double *At = tls_scratch((size_t)m * k);                 // A^T, column-major m x k
  for (int i = 0; i < m; i++)
    for (int l = 0; l < k; l++) At[i + (size_t)l * m] = A.a[l + (size_t)i * k];
  gemm_core(m, n, k, alpha, At, m, B.a.data(), 1, k, beta, C.a.data(), m);
}
/// C = beta*C + alpha * A * B^T
inline void gemm_nt(double alpha, const Mat &A, const Mat &B, double beta, Mat &C) {
  const int m = A.r, k = A.c, n = B.r;
  assert(B.c == k && C.r == m && C.c == n);
  if (m == 0 || n == 0) return;
  gemm_core(m, n, k, alpha, A.a.data(), m, B.a.data(), n, 1, beta, C.a.data(), m);
}
/// C = beta*C + alpha * A^T * B^T
inline void gemm_tt(double alpha, const Mat &A, const Mat &B, double beta, Mat &C) {
  const int m = A.c, k = A.r, n = B.r;
  assert(B.c == k && C.r == m && C.c == n);
  if (m == 0 || n == 0) return;
  double *At = tls_scratch((size_t)m * k);
  for (int i = 0; i < m; i++)
    for (int l = 0; l < k; l++) At[i + (size_t)l * m] = A.a[l + (size_t)i * k];
  gemm_core(m, n, k, alpha, At, m, B.a.data(), n, 1, beta, C.a.data(), m);
}

inline Mat transpose(const Mat &A) {
  Mat T(A.c, A.r);
  for (int j = 0; j < A.c; j++)
    for (int i = 0; i < A.r; i++) T(j, i) = A(i, j);
  return T;
}
inline void axpy(double alpha, const Mat &X, Mat &Y) {
  assert(X.a.size() == Y.a.size());
  for (size_t i = 0; i < X.a.size(); i++) Y.a[i] += alpha * X.a[i];
}
inline void scale(Mat &X, double s) { for (auto &v : X.a) v *= s; }
inline double inf_norm(const Mat &X) {
  double m = 0.0;
  for (double v : X.a) { double f = std::fabs(v); if (!(f <= m)) m = f; } // NaN propagates
  return m;
}
/// A = A.selfadjointView<Lower>()  (mirror the lower triangle; riccati-kernel.hxx:180,190,206)
inline void symmetrize_from_lower(Mat &A) {
  assert(A.r == A.c);
  for (int j = 0; j < A.c; j++)
    for (int i = j + 1; i < A.r; i++) A(j, i) = A(i, j);
}

// ---------------------------------------------------------------------------------
// PartialPivLU (Eigen 3.4.0 semantics == LAPACK dgetf2): P A = L U.
// ---------------------------------------------------------------------------------
struct PartialPivLU {
  Mat lu;
  std::vector<int> piv; // row transpositions: at step k rows k and piv[k] were swapped
  bool ok = true;
  void compute(const Mat &A) {
    const int n = A.r;
    assert(A.c == n);
    lu = A;
    piv.assign(n, 0);
    ok = true;
    for (int k = 0; k < n; k++) {
      int p = k;
      double big = std::fabs(lu(k, k));
      for (int i = k + 1; i < n; i++) {
        double v = std::fabs(lu(i, k));
        if (v > big) { big = v; p = i; }
      }
      piv[k] = p;
      if (big == 0.0) { ok = false; continue; }
      if (p != k)
        for (int j = 0; j < n; j++) std::swap(lu(k, j), lu(p, j));
      const double d = lu(k, k);
      for (int i = k + 1; i < n; i++) lu(i, k) /= d;
      for (int j = k + 1; j < n; j++) {
        const double u = lu(k, j);
        double *cj = lu.col(j);
        const double *lk = lu.col(k);
        for (int i = k + 1; i < n; i++) cj[i] -= lk[i] * u;
      }
    }
  }
  /// X <- A^{-1} X
  void solveInPlace(Mat &X) const {
    const int n = lu.r;
    assert(X.r == n);
    for (int k = 0; k < n; k++)
      if (piv[k] != k)
        for (int j = 0; j < X.c; j++) std::swap(X(k, j), X(piv[k], j));
    // four right-hand sides per pass: every column of L / U is loaded once for four columns of X (the
    // per-column operation order is unchanged, so are the bits)
    const int nc4 = X.c & ~3;
    for (int j = 0; j < nc4; j += 4) {
      double *x0 = X.col(j), *x1 = X.col(j + 1), *x2 = X.col(j + 2), *x3 = X.col(j + 3);
      for (int k = 0; k < n; k++) {           // unit-lower forward substitution
        const double a0 = x0[k], a1 = x1[k], a2 = x2[k], a3 = x3[k];
        const double *lk = lu.col(k);
#pragma omp simd
        for (int i = k + 1; i < n; i++) { const double l = lk[i]; x0[i] -= l * a0; x1[i] -= l * a1; x2[i] -= l * a2; x3[i] -= l * a3; }
      }
      for (int k = n - 1; k >= 0; k--) {      // upper back substitution
        const double dk = lu(k, k);
        x0[k] /= dk; x1[k] /= dk; x2[k] /= dk; x3[k] /= dk;
        const double a0 = x0[k], a1 = x1[k], a2 = x2[k], a3 = x3[k];
        const double *uk = lu.col(k);
#pragma omp simd
        for (int i = 0; i < k; i++) { const double u = uk[i]; x0[i] -= u * a0; x1[i] -= u * a1; x2[i] -= u * a2; x3[i] -= u * a3; }
      }
    }
    for (int j = nc4; j < X.c; j++) {
      double *x = X.col(j);
      for (int k = 0; k < n; k++) {           // unit-lower forward substitution
        const double xk = x[k];
        const double *lk = lu.col(k);
        for (int i = k + 1; i < n; i++) x[i] -= lk[i] * xk;
      }
      for (int k = n - 1; k >= 0; k--) {      // upper back substitution
        x[k] /= lu(k, k);
        const double xk = x[k];
        const double *uk = lu.col(k);
        for (int i = 0; i < k; i++) x[i] -= uk[i] * xk;
      }
    }
  }
};

// ---------------------------------------------------------------------------------
// LLT (Eigen 3.4.0 semantics == LAPACK dpotf2, lower): A = L L^T.
// ---------------------------------------------------------------------------------
struct LLT {
  Mat l;
  bool ok = true;
  void compute(const Mat &A) {
    const int n = A.r;
    l = A;
    ok = true;
    for (int k = 0; k < n; k++) {
      double d = l(k, k);
      for (int p = 0; p < k; p++) d -= l(k, p) * l(k, p);
      if (!(d > 0.0)) { ok = false; d = std::fabs(d) + 1e-300; }
      d = std::sqrt(d);
      l(k, k) = d;
      for (int i = k + 1; i < n; i++) {
        double s = l(i, k);
        for (int p = 0; p < k; p++) s -= l(i, p) * l(k, p);
        l(i, k) = s / d;
      }
    }
  }
  /// X <- A^{-1} X
  void solveInPlace(Mat &X) const {
    const int n = l.r;
    // four right-hand sides per pass (same per-column operation order as the one-column loop below)
    const int nc4 = X.c & ~3;
    for (int j = 0; j < nc4; j += 4) {
      double *x0 = X.col(j), *x1 = X.col(j + 1), *x2 = X.col(j + 2), *x3 = X.col(j + 3);
      for (int k = 0; k < n; k++) {
        const double dk = l(k, k);
        x0[k] /= dk; x1[k] /= dk; x2[k] /= dk; x3[k] /= dk;
        const double a0 = x0[k], a1 = x1[k], a2 = x2[k], a3 = x3[k];
        const double *lk = l.col(k);
#pragma omp simd
        for (int i = k + 1; i < n; i++) { const double v = lk[i]; x0[i] -= v * a0; x1[i] -= v * a1; x2[i] -= v * a2; x3[i] -= v * a3; }
      }
      for (int k = n - 1; k >= 0; k--) {
        double s0 = x0[k], s1 = x1[k], s2 = x2[k], s3 = x3[k];
        const double *lk = l.col(k);
        for (int i = k + 1; i < n; i++) { const double v = lk[i]; s0 -= v * x0[i]; s1 -= v * x1[i]; s2 -= v * x2[i]; s3 -= v * x3[i]; }
        const double dk = l(k, k);
        x0[k] = s0 / dk; x1[k] = s1 / dk; x2[k] = s2 / dk; x3[k] = s3 / dk;
      }
    }
    for (int j = nc4; j < X.c; j++) {
      double *x = X.col(j);
      for (int k = 0; k < n; k++) {
        x[k] /= l(k, k);
        const double xk = x[k];
        const double *lk = l.col(k);
        for (int i = k + 1; i < n; i++) x[i] -= lk[i] * xk;
      }
      for (int k = n - 1; k >= 0; k--) {
        double s = x[k];
        const double *lk = l.col(k);
        for (int i = k + 1; i < n; i++) s -= lk[i] * x[i];
        x[k] = s / l(k, k);
      }
    }
  }
};

} // namespace gar_oracle
