// oracle/bunchkaufman.hpp — TEST INFRASTRUCTURE ONLY.
//
// Restatement of the reference's symmetric-indefinite LDL^T with Bunch-Kaufman partial
// pivoting:  P A P^T = L D L^T, lower storage, 1x1 / 2x2 pivots, INVERTED diagonal
// blocks stored in place.
//   factor : /root/reference/include/aligator/core/bunchkaufman.hpp:20-167 (unblocked)
//   post   : :389-415  (sub-diagonal of the 2x2 blocks split out, rows of the already
//                       computed L columns swapped so that L is a true unit-lower factor
//                       of the fully permuted matrix)
//   solve  : :447-516
// The reference switches to a blocked (block 32) variant for n > 32 (:360-367, :529);
// it selects pivots with the same rule on the same updated columns and differs from
// the unblocked one only in summation order, so the unblocked algorithm is restated
// for every size.
#pragma once
#include "linalg.hpp"

namespace gar_oracle {

struct BunchKaufman {
  Mat m;                    // L (strict lower, unit diagonal implied) + inverted D blocks
  std::vector<double> sub;  // sub-diagonal of the inverted 2x2 blocks
  std::vector<int> piv;     // p >= 0: 1x1 pivot, row k <-> p ; p < 0: 2x2, row k+1 <-> -1-p
  bool ok = true;

  static double alpha() { return (1.0 + std::sqrt(17.0)) / 8.0; }

  /// Only the LOWER triangle of A is read (bunchkaufman.hpp:667-668).
  void compute(const Mat &A) {
    const int n = A.r;
    assert(A.c == n);
    m.resize(n, n);
    for (int j = 0; j < n; j++)
      for (int i = j; i < n; i++) m(i, j) = A(i, j);
    sub.assign(n, 0.0);
    piv.assign(n, 0);
    ok = factor();
    if (!ok) return;
    // split off the sub-diagonal of the 2x2 blocks (bunchkaufman.hpp:391-402)
    for (int k = 0; k < n;) {
      if (piv[k] < 0) {
        sub[k] = m(k + 1, k);
        sub[k + 1] = 0.0;
        m(k + 1, k) = 0.0;
        k += 2;
      } else {
        sub[k] = 0.0;
        k += 1;
      }
    }
    // apply the later interchanges to the earlier columns of L (:404-415)
    for (int k = 0; k < n;) {
      int p = piv[k];
      if (p < 0) {
        p = -1 - p;
        for (int j = 0; j < k; j++) std::swap(m(k + 1, j), m(p, j));
        k += 2;
      } else {
        for (int j = 0; j < k; j++) std::swap(m(k, j), m(p, j));
        k += 1;
      }
    }
  }

  /// X <- A^{-1} X   (bunchkaufman.hpp:447-516)
  void solveInPlace(Mat &X) const {
    const int n = m.r;
    const int nr = X.c;
    assert(X.r == n);
    auto swap_rows = [&](int a, int b) {
      if (a == b) return;
      for (int j = 0; j < nr; j++) std::swap(X(a, j), X(b, j));
    };
    for (int k = 0; k < n;) {
      int p = piv[k];
      if (p < 0) { swap_rows(k + 1, -1 - p); k += 2; }
      else { swap_rows(k, p); k += 1; }
    }
    for (int j = 0; j < nr; j++) {  // unit-lower solve
      double *x = X.col(j);
      for (int k = 0; k < n; k++) {
        const double xk = x[k];
        const double *lk = m.col(k);
        for (int i = k + 1; i < n; i++) x[i] -= lk[i] * xk;
      }
    }
    for (int k = 0; k < n;) {       // block-diagonal solve with the stored inverses
      if (piv[k] < 0) {
        const double akp1k = sub[k], ak = m(k, k), akp1 = m(k + 1, k + 1);
        for (int j = 0; j < nr; j++) {
          const double xk = X(k, j), xkp1 = X(k + 1, j);
          X(k, j) = xk * ak + xkp1 * akp1k;
          X(k + 1, j) = xkp1 * akp1 + xk * akp1k;
        }
        k += 2;
      } else {
        const double dk = m(k, k);
        for (int j = 0; j < nr; j++) X(k, j) *= dk;
        k += 1;
      }
    }
    for (int j = 0; j < nr; j++) {  // unit-upper (L^T) solve
      double *x = X.col(j);
      for (int k = n - 1; k >= 0; k--) {
        double s = x[k];
        const double *lk = m.col(k);
        for (int i = k + 1; i < n; i++) s -= lk[i] * x[i];
        x[k] = s;
      }
    }
    for (int k = n; k > 0;) {       // undo the interchanges in reverse order
      k -= 1;
      int p = piv[k];
      if (p < 0) { swap_rows(k, -1 - p); k -= 1; }
      else swap_rows(k, p);
    }
  }

private:
  bool factor() {
    const int n = m.r;
    const double al = alpha();
    if (n == 0) return true;
    if (n == 1) {
      if (std::fabs(m(0, 0)) == 0.0) return false;
      m(0, 0) = 1.0 / m(0, 0);
      return true;
    }
    int k = 0;
    while (k < n) {
      int kstep = 1;
      const double abs_akk = std::fabs(m(k, k));
      int imax = k + 1;
      double colmax = 0.0;
      for (int i = k + 1; i < n; i++) {  // first maximum, like Eigen's maxCoeff(&idx)
        const double v = std::fabs(m(i, k));
        if (v > colmax) { colmax = v; imax = i; }
      }
      if (k + 1 >= n) imax = k + 1;
      int kp;
      if (std::max(abs_akk, colmax) == 0.0) return false;
      if (abs_akk >= colmax * al) {
        kp = k;
      } else {
        double rowmax = 0.0;
        for (int j = k; j < imax; j++) rowmax = std::max(rowmax, std::fabs(m(imax, j)));
        for (int i = imax + 1; i < n; i++) rowmax = std::max(rowmax, std::fabs(m(i, imax)));
        if (abs_akk >= (al * colmax) * (colmax / rowmax)) {
          kp = k;
        } else if (std::fabs(m(imax, imax)) >= al * rowmax) {
          kp = imax;
        } else {
          kp = imax;
          kstep = 2;
        }
      }
      const int kk = k + kstep - 1;
      if (kp != kk) {  // symmetric interchange of rows/cols kk and kp in the trailing block
        for (int i = kp + 1; i < n; i++) std::swap(m(i, kk), m(i, kp));
        for (int j = kk + 1; j < kp; j++) std::swap(m(j, kk), m(kp, j));
        std::swap(m(kk, kk), m(kp, kp));
        if (kstep == 2) std::swap(m(k + 1, k), m(kp, k));
      }
      if (kstep == 1) {
        const double d11 = 1.0 / m(k, k);
        m(k, k) = d11;
        for (int j = k + 1; j < n; j++) {
          const double d11xj = m(j, k) * d11;
          for (int i = j; i < n; i++) m(i, j) -= d11xj * m(i, k);
        }
        for (int i = k + 1; i < n; i++) m(i, k) *= d11;
      } else {
        const double d21_abs = std::fabs(m(k + 1, k));
        const double d21_inv = 1.0 / d21_abs;
        const double d11 = d21_inv * m(k + 1, k + 1);
        const double d22 = d21_inv * m(k, k);
        const double t = 1.0 / ((d11 * d22) - 1.0);
        const double d = t * d21_inv;
        const double d21 = m(k + 1, k) * d21_inv;
        m(k, k) = d11 * d;
        m(k + 1, k) = -d21 * d;
        m(k + 1, k + 1) = d22 * d;
        for (int j = k + 2; j < n; j++) {
          const double wk = ((m(j, k) * d11) - (m(j, k + 1) * d21)) * d;
          const double wkp1 = ((m(j, k + 1) * d22) - (m(j, k) * d21)) * d;
          for (int i = j; i < n; i++) m(i, j) -= m(i, k) * wk + m(i, k + 1) * wkp1;
          m(j, k) = wk;
          m(j, k + 1) = wkp1;
        }
      }
      if (kstep == 1) piv[k] = kp;
      else { piv[k] = -1 - kp; piv[k + 1] = -1 - kp; }
      k += kstep;
    }
    return true;
  }
};

} // namespace gar_oracle
