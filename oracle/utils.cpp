// oracle/utils.cpp — TEST INFRASTRUCTURE ONLY.
// Restatement of /root/reference/include/aligator/gar/utils.{hpp,hxx}:
// lqrInitializeSolution, lqrComputeKktError, lqrNumRows, lqrDenseMatrix.
#include "gar_oracle.hpp"

namespace gar_oracle {

// utils.hpp:139-168
void lqrInitializeSolution(const Problem &p, std::vector<Mat> &xs, std::vector<Mat> &us,
                           std::vector<Mat> &vs, std::vector<Mat> &lbdas) {
  const int N = p.horizon();
  xs.assign(N + 1, Mat());
  us.assign(N + 1, Mat());
  vs.assign(N + 1, Mat());
  lbdas.assign(N + 1, Mat());
  lbdas[0].resize(p.nc0());
  for (int i = 0; i <= N; i++) {
    const Knot &kn = p.stages[i];
    xs[i].resize(kn.nx);
    us[i].resize(kn.nu);
    vs[i].resize(kn.nc);
    if (i == N) break;
    lbdas[i + 1].resize(kn.nx2);
  }
  if (p.stages.back().nu == 0) us.pop_back();
}

// utils.hxx:94-192
std::array<double, 3> lqrComputeKktError(const Problem &p, const std::vector<Mat> &xs,
                                         const std::vector<Mat> &us, const std::vector<Mat> &vs,
                                         const std::vector<Mat> &lbdas, double mudyn,
                                         double mueq, const Mat *theta) {
  const int N = p.horizon();
  double dynErr = 0.0, cstErr = 0.0, dualErr = 0.0;
  {
    Mat dyn = p.g0;                                     // :122-126
    gemm_nn(1.0, p.G0, xs[0], 1.0, dyn);
    axpy(-mudyn, lbdas[0], dyn);
    dynErr = std::max(dynErr, inf_norm(dyn));
  }
  for (int t = 0; t <= N; t++) {
    const Knot &k = p.stages[t];
    Mat gx(k.nx), gu(k.nu);
    Mat cst = k.d;                                      // :143
    gemm_nn(1.0, k.C, xs[t], 1.0, cst);
    axpy(-mueq, vs[t], cst);
    gx = k.q;                                           // :144
    gemm_nn(1.0, k.Q, xs[t], 1.0, gx);
    gemm_tn(1.0, k.C, vs[t], 1.0, gx);
    gu = k.r;                                           // :145
    gemm_tn(1.0, k.S, xs[t], 1.0, gu);
    gemm_tn(1.0, k.D, vs[t], 1.0, gu);
    if (k.nu > 0) {                                     // :147-151
      gemm_nn(1.0, k.D, us[t], 1.0, cst);
      gemm_nn(1.0, k.S, us[t], 1.0, gx);
      gemm_nn(1.0, k.R, us[t], 1.0, gu);
    }
    if (t == 0) gemm_tn(1.0, p.G0, lbdas[0], 1.0, gx);  // :153-158
    else gemm_tn(1.0, p.stages[t - 1].E, lbdas[t], 1.0, gx);
    if (t < N) {                                        // :160-171
      Mat dyn = k.f;
      gemm_nn(1.0, k.A, xs[t], 1.0, dyn);
      gemm_nn(1.0, k.B, us[t], 1.0, dyn);
      gemm_nn(1.0, k.E, xs[t + 1], 1.0, dyn);
      axpy(-mudyn, lbdas[t + 1], dyn);
      gemm_tn(1.0, k.A, lbdas[t + 1], 1.0, gx);
      gemm_tn(1.0, k.B, lbdas[t + 1], 1.0, gu);
      dynErr = std::max(dynErr, inf_norm(dyn));
    }
    if (theta && k.nth > 0) {                           // :173-182 (the _gt part is unused)
      gemm_nn(1.0, k.Gx, *theta, 1.0, gx);
      gemm_nn(1.0, k.Gu, *theta, 1.0, gu);
    }
    dualErr = std::max({dualErr, inf_norm(gx), inf_norm(gu)});
    cstErr = std::max(cstErr, inf_norm(cst));
  }
  return {dynErr, cstErr, dualErr};
}

// utils.hpp:71-84
int lqrNumRows(const Problem &p) {
  const int N = p.horizon();
  int nrows = p.nc0();
  for (int t = 0; t <= N; t++) {
    const Knot &m = p.stages[t];
    nrows += m.nx + m.nu + m.nc;
    if (t != N) nrows += m.nx;
  }
  return nrows;
}

// utils.hxx:194-267.  Unknown ordering [lbd0 | x0 u0 v0 lbd1 | x1 u1 v1 lbd2 | ...]
// (utils.hpp:106-137); the system is  mat * z = -rhs.
bool lqrDenseMatrix(const Problem &p, double mudyn, double mueq, Mat &mat, Mat &rhs) {
  if (p.stages.empty()) return false;
  const int N = p.horizon();
  const int nrows = lqrNumRows(p);
  mat.resize(nrows, nrows);
  rhs.resize(nrows);
  int idx = 0;
  {
    const int nc0 = p.nc0(), nx0 = p.stages[0].nx;
    for (int j = 0; j < nx0; j++)
      for (int i = 0; i < nc0; i++) { mat(i, nc0 + j) = p.G0(i, j); mat(nc0 + j, i) = p.G0(i, j); }
    for (int i = 0; i < nc0; i++) { mat(i, i) = -mudyn; rhs[i] = p.g0[i]; }
    idx += nc0;
  }
  for (int t = 0; t <= N; t++) {
    const Knot &m = p.stages[t];
    const int nx = m.nx, nu = m.nu, nc = m.nc, n = nx + nu + nc;
    // lower triangle of the (x,u,v) block, then mirrored (:239-245)
    for (int j = 0; j < nx; j++)
      for (int i = j; i < nx; i++) mat(idx + i, idx + j) = m.Q(i, j);
    for (int j = 0; j < nx; j++)
      for (int i = 0; i < nu; i++) mat(idx + nx + i, idx + j) = m.S(j, i);
    for (int j = 0; j < nu; j++)
      for (int i = j; i < nu; i++) mat(idx + nx + i, idx + nx + j) = m.R(i, j);
    for (int j = 0; j < nx; j++)
      for (int i = 0; i < nc; i++) mat(idx + nx + nu + i, idx + j) = m.C(i, j);
    for (int j = 0; j < nu; j++)
      for (int i = 0; i < nc; i++) mat(idx + nx + nu + i, idx + nx + j) = m.D(i, j);
    for (int i = 0; i < nc; i++) mat(idx + nx + nu + i, idx + nx + nu + i) = -mueq;
    for (int j = 0; j < n; j++)
      for (int i = j + 1; i < n; i++) mat(idx + j, idx + i) = mat(idx + i, idx + j);
    for (int i = 0; i < nx; i++) rhs[idx + i] = m.q[i];
    for (int i = 0; i < nu; i++) rhs[idx + nx + i] = m.r[i];
    for (int i = 0; i < nc; i++) rhs[idx + nx + nu + i] = m.d[i];
    if (t != N) {                                       // dynamics row [A B 0 -mu I E]
      const int r0 = idx + n;
      for (int i = 0; i < nx; i++) {
        for (int j = 0; j < nx; j++) { mat(r0 + i, idx + j) = m.A(i, j); mat(idx + j, r0 + i) = m.A(i, j); }
        for (int j = 0; j < nu; j++) { mat(r0 + i, idx + nx + j) = m.B(i, j); mat(idx + nx + j, r0 + i) = m.B(i, j); }
        mat(r0 + i, r0 + i) = -mudyn;
        for (int j = 0; j < nx; j++) { mat(r0 + i, r0 + nx + j) = m.E(i, j); mat(r0 + nx + j, r0 + i) = m.E(i, j); }
        rhs[r0 + i] = m.f[i];
      }
      idx += n + nx;
    }
  }
  return true;
}

} // namespace gar_oracle
