// oracle/dense.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's dense per-knot KKT
// solver:
//   DenseKernel::terminalSolve / stageKernelSolve / forwardStep   dense-kernel.hpp:56-95, 97-172, 174-212
//   RiccatiSolverDense::backward / forward                        dense-riccati.hxx:46-118
// One Bunch-Kaufman factorization of the (nu + nc + 2 nx)^2 stage KKT matrix
//   [ R   D^T   B^T    0   ]        unknown block rows [K; Z; L; Y] — the same row layout as the proximal
//   [ D  -mueq  0      0   ]        solver's [K; Z; L; A] gains, so StageFactor::{ff, fb, fth} and
//   [ B   0    -mudyn  E   ]        forwardImpl are shared with riccati.cpp.
//   [ 0   0     E^T    Pxx+]
// Unlike the proximal kernel it READS Gv (dense-kernel.hpp:76,134).
#include "gar_oracle.hpp"

namespace gar_oracle {

static void neg_into(Mat &dst, int r0, int c0, const Mat &src, bool transpose = false) {
  if (!transpose) { for (int j = 0; j < src.c; j++) for (int i = 0; i < src.r; i++) dst(r0 + i, c0 + j) = -src(i, j); }
  else { for (int j = 0; j < src.c; j++) for (int i = 0; i < src.r; i++) dst(r0 + j, c0 + i) = -src(i, j); }
}
static void put_into(Mat &dst, int r0, int c0, const Mat &src, bool transpose = false) {
  if (!transpose) { for (int j = 0; j < src.c; j++) for (int i = 0; i < src.r; i++) dst(r0 + i, c0 + j) = src(i, j); }
  else { for (int j = 0; j < src.c; j++) for (int i = 0; i < src.r; i++) dst(r0 + j, c0 + i) = src(i, j); }
}
static Mat rows_of(const Mat &M, int r0, int nr) {
  Mat out(nr, M.c);
  for (int j = 0; j < M.c; j++) for (int i = 0; i < nr; i++) out(i, j) = M(r0 + i, j);
  return out;
}

// dense-kernel.hpp:56-95 (nt == false) and :97-172; vn == nullptr at the terminal knot
static void dense_stage(const Knot &k, StageFactor &d, const Value *vn, bool terminal, double mudyn, double mueq) {
  const int nx = k.nx, nu = k.nu, nc = k.nc, nth = k.nth;
  const int nr = terminal ? nu + nc : nu + nc + 2 * nx;
  const int oZ = nu, oL = nu + nc, oY = nu + nc + nx;
  Mat kkt(nr, nr);
  put_into(kkt, 0, 0, k.R);
  put_into(kkt, oZ, 0, k.D);
  put_into(kkt, 0, oZ, k.D, true);
  for (int i = 0; i < nc; i++) kkt(oZ + i, oZ + i) = -mueq;
  if (!terminal) {
    put_into(kkt, oL, 0, k.B);
    put_into(kkt, 0, oL, k.B, true);
    for (int i = 0; i < nx; i++) kkt(oL + i, oL + i) = -mudyn;
    put_into(kkt, oL, oY, k.E);
    put_into(kkt, oY, oL, k.E, true);
    if (vn) put_into(kkt, oY, oY, vn->Pmat);
  }
  BunchKaufman ldl;
  ldl.compute(kkt);                                    // 1. factorize
  Mat ff(nr), fb(nr, nx), ft(nr, nth);                 // 2. right-hand sides
  neg_into(ff, 0, 0, k.r); neg_into(ff, oZ, 0, k.d);
  neg_into(fb, 0, 0, k.S, true); neg_into(fb, oZ, 0, k.C);
  neg_into(ft, 0, 0, k.Gu); neg_into(ft, oZ, 0, k.Gv);
  if (!terminal) {
    neg_into(ff, oL, 0, k.f);
    if (vn) neg_into(ff, oY, 0, vn->pvec);
    neg_into(fb, oL, 0, k.A);
    if (vn) neg_into(ft, oY, 0, vn->Vxt);
  }
  if (nr > 0) { ldl.solveInPlace(ff); ldl.solveInPlace(fb); if (nth > 0) ldl.solveInPlace(ft); }
  const Mat kff = rows_of(ff, 0, nu), zff = rows_of(ff, oZ, nc), K = rows_of(fb, 0, nu), Z = rows_of(fb, oZ, nc);
  const Mat Kth = rows_of(ft, 0, nu), Zth = rows_of(ft, oZ, nc);
  Value &v = d.vm;                                     // 3. value function
  v.Pmat = k.Q;  gemm_nn(1.0, k.S, K, 1.0, v.Pmat);  gemm_tn(1.0, k.C, Z, 1.0, v.Pmat);
  v.pvec = k.q;  gemm_nn(1.0, k.S, kff, 1.0, v.pvec); gemm_tn(1.0, k.C, zff, 1.0, v.pvec);
  if (nth > 0) {
    v.Vxt = k.Gx;    gemm_tn(1.0, K, k.Gu, 1.0, v.Vxt);     gemm_tn(1.0, Z, k.Gv, 1.0, v.Vxt);
    v.Vtt = k.Gth;   gemm_tn(1.0, Kth, k.Gu, 1.0, v.Vtt);   gemm_tn(1.0, Zth, k.Gv, 1.0, v.Vtt);
    v.vt = k.gamma;  gemm_tn(1.0, k.Gu, kff, 1.0, v.vt);    gemm_tn(1.0, k.Gv, zff, 1.0, v.vt);
  }
  if (!terminal) {
    const Mat lff = rows_of(ff, oL, nx), L = rows_of(fb, oL, nx);
    gemm_tn(1.0, k.A, L, 1.0, v.Pmat);
    gemm_tn(1.0, k.A, lff, 1.0, v.pvec);
    if (nth > 0 && vn) {
      const Mat Y = rows_of(fb, oY, nx), Yth = rows_of(ft, oY, nx), yff = rows_of(ff, oY, nx);
      gemm_tn(1.0, Y, vn->Vxt, 1.0, v.Vxt);
      gemm_tn(1.0, Yth, vn->Vxt, 1.0, v.Vtt);
      gemm_tn(1.0, vn->Vxt, yff, 1.0, v.vt);
    }
  }
  d.ff.setZero(); d.fb.setZero(); d.fth.setZero();     // block rows [K; Z; L; Y] (the terminal knot has none of L, Y)
  for (int i = 0; i < nr; i++) {
    d.ff[i] = ff[i];
    for (int j = 0; j < nx; j++) d.fb(i, j) = fb(i, j);
    for (int j = 0; j < nth; j++) d.fth(i, j) = ft(i, j);
  }
}

RiccatiSolverDense::RiccatiSolverDense(const Problem &p) : problem(&p) {
  for (const Knot &k : p.stages) datas.emplace_back(k.nx, k.nu, k.nc, k.nth);
  const int nth = p.ntheta();
  thGrad.resize(nth);
  thHess.resize(nth, nth);
}

// dense-riccati.hxx:46-94
bool RiccatiSolverDense::backward(double mudyn, double mueq) {
  const Problem &p = *problem;
  const int N = p.horizon();
  dense_stage(p.stages[N], datas[N], nullptr, true, mudyn, mueq);
  for (int i = N - 1; i >= 0; i--) dense_stage(p.stages[i], datas[i], &datas[i + 1].vm, false, mudyn, mueq);
  const Value &v0 = datas[0].vm;                      // initial stage (:68-92)
  const int nx = p.stages[0].nx, nc0 = p.nc0(), nth = p.ntheta();
  Mat kkt(nx + nc0, nx + nc0);
  put_into(kkt, 0, 0, v0.Pmat);
  put_into(kkt, nx, 0, p.G0);
  put_into(kkt, 0, nx, p.G0, true);
  for (int i = 0; i < nc0; i++) kkt(nx + i, nx + i) = -mudyn;
  kkt0.chol.compute(kkt);
  kkt0.ff.resize(nx + nc0);
  for (int i = 0; i < nx; i++) kkt0.ff[i] = -v0.pvec[i];
  for (int i = 0; i < nc0; i++) kkt0.ff[nx + i] = -p.g0[i];
  kkt0.chol.solveInPlace(kkt0.ff);
  kkt0.fth.resize(nx + nc0, nth);
  for (int j = 0; j < nth; j++) for (int i = 0; i < nx; i++) kkt0.fth(i, j) = -v0.Vxt(i, j);
  if (nth > 0) kkt0.chol.solveInPlace(kkt0.fth);
  thGrad = v0.vt;
  thHess = v0.Vtt;
  for (int a = 0; a < nth; a++) {
    double s = 0.0;
    for (int i = 0; i < nx; i++) s += v0.Vxt(i, a) * kkt0.ff[i];
    thGrad[a] += s;
    for (int b = 0; b < nth; b++) {
      double h = 0.0;
      for (int i = 0; i < nx; i++) h += v0.Vxt(i, a) * kkt0.fth(i, b);
      thHess(a, b) += h;
    }
  }
  return true;
}

// dense-riccati.hxx:96-118 + DenseKernel::forwardStep (dense-kernel.hpp:174-212): the same recursion as
// riccati-kernel.hxx:303-361 on the same block rows
bool RiccatiSolverDense::forward(std::vector<Mat> &xs, std::vector<Mat> &us, std::vector<Mat> &vs,
                                 std::vector<Mat> &lbdas, const Mat *theta) const {
  const Problem &p = *problem;
  const int nx = p.stages[0].nx, nc0 = p.nc0(), nth = p.ntheta();
  for (int i = 0; i < nx; i++) xs[0][i] = kkt0.ff[i];
  for (int i = 0; i < nc0; i++) lbdas[0][i] = kkt0.ff[nx + i];
  if (theta)
    for (int j = 0; j < nth; j++) {
      for (int i = 0; i < nx; i++) xs[0][i] += kkt0.fth(i, j) * (*theta)[j];
      for (int i = 0; i < nc0; i++) lbdas[0][i] += kkt0.fth(nx + i, j) * (*theta)[j];
    }
  return forwardImpl(p.stages, datas, 0, (int)p.stages.size(), xs, us, vs, lbdas, theta);
}

} // namespace gar_oracle
