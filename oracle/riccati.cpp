// oracle/riccati.cpp — TEST INFRASTRUCTURE ONLY.
// Restatement of ProximalRiccatiKernel / ProximalRiccatiSolver
// (/root/reference/include/aligator/gar/riccati-kernel.hxx, proximal-riccati.hxx).
#include "gar_oracle.hpp"

namespace gar_oracle {

Knot::Knot(int nx_, int nu_, int nc_, int nth_)
    : nx(nx_), nu(nu_), nc(nc_), nx2(nx_), nth(nth_), Q(nx_, nx_), S(nx_, nu_), R(nu_, nu_),
      q(nx_), r(nu_), A(nx_, nx_), B(nx_, nu_), E(nx_, nx_), f(nx_), C(nc_, nx_), D(nc_, nu_),
      d(nc_), Gth(nth_, nth_), Gx(nx_, nth_), Gu(nu_, nth_), Gv(nc_, nth_), gamma(nth_) {}

void Knot::addParameterization(int nth_) {
  nth = nth_;
  Gth.resize(nth, nth);
  Gx.resize(nx, nth);
  Gu.resize(nu, nth);
  Gv.resize(nc, nth);
  gamma.resize(nth);
}

Value::Value(int nx, int nth)
    : Pmat(nx, nx), pvec(nx), Vxx(nx, nx), vx(nx), Vxt(nx, nth), Vtt(nth, nth), vt(nth) {}

StageFactor::StageFactor(int nx_, int nu_, int nc_, int nth_)
    : nx(nx_), nu(nu_), nc(nc_), nth(nth_), ff(nu_ + nc_ + 2 * nx_), fb(nu_ + nc_ + 2 * nx_, nx_),
      fth(nu_ + nc_ + 2 * nx_, nth_), vm(nx_, nth_), Qhat(nx_, nx_), Rhat(nu_, nu_),
      Shat(nx_, nu_), qhat(nx_), rhat(nu_), AtV(nx_, nx_), BtV(nu_, nx_), Guhat(nu_, nth_),
      Einv(nx_, nx_), Ptilde(nx_, nx_) {}

namespace {
// copy `src` into rows [r0, r0+src.r) of dst (all columns)
void put_rows(Mat &dst, int r0, const Mat &src) {
  assert(dst.c == src.c);
  for (int j = 0; j < src.c; j++)
    for (int i = 0; i < src.r; i++) dst(r0 + i, j) = src(i, j);
}
Mat get_rows(const Mat &src, int r0, int nr) {
  Mat out(nr, src.c);
  for (int j = 0; j < src.c; j++)
    for (int i = 0; i < nr; i++) out(i, j) = src(r0 + i, j);
  return out;
}

// assemble [Rm D^T; D -mueq I] (lower triangle is what BunchKaufman reads)
Mat assemble_reduced_kkt(const Mat &Rm, const Mat &D, double mueq) {
  const int nu = Rm.r, nc = D.r;
  Mat K(nu + nc, nu + nc);
  for (int j = 0; j < nu; j++)
    for (int i = 0; i < nu; i++) K(i, j) = Rm(i, j);
  for (int j = 0; j < nu; j++)
    for (int i = 0; i < nc; i++) { K(nu + i, j) = D(i, j); K(j, nu + i) = D(i, j); }
  for (int i = 0; i < nc; i++) K(nu + i, nu + i) = -mueq;
  return K;
}
} // namespace

// riccati-kernel.hxx:90-150
void terminalSolve(const Knot &model, double mueq, StageFactor &d) {
  const int nx = model.nx, nu = model.nu, nc = model.nc, nth = model.nth;
  Value &vc = d.vm;
  Mat kff(nu), zff(nc), K(nu, nx), Z(nc, nx), Kth(nu, nth), Zth(nc, nth);

  if (nu == 0) {                                   // :105-108
    for (int j = 0; j < nx; j++)
      for (int i = 0; i < nc; i++) Z(i, j) = model.C(i, j) / mueq;
    for (int i = 0; i < nc; i++) zff[i] = model.d[i] / mueq;
  } else {                                         // :110-131
    Mat kkt = assemble_reduced_kkt(model.R, model.D, mueq);
    d.kktChol.compute(kkt);
    Mat rhs(nu + nc, 1 + nx + nth);                // columns [ff | fb | fth]
    for (int i = 0; i < nu; i++) rhs(i, 0) = -model.r[i];
    for (int i = 0; i < nc; i++) rhs(nu + i, 0) = -model.d[i];
    for (int j = 0; j < nx; j++) {
      for (int i = 0; i < nu; i++) rhs(i, 1 + j) = -model.S(j, i);
      for (int i = 0; i < nc; i++) rhs(nu + i, 1 + j) = -model.C(i, j);
    }
    for (int j = 0; j < nth; j++)
      for (int i = 0; i < nu; i++) rhs(i, 1 + nx + j) = -model.Gu(i, j);  // Zth rows = 0
    d.kktChol.solveInPlace(rhs);
    for (int i = 0; i < nu; i++) kff[i] = rhs(i, 0);
    for (int i = 0; i < nc; i++) zff[i] = rhs(nu + i, 0);
    for (int j = 0; j < nx; j++) {
      for (int i = 0; i < nu; i++) K(i, j) = rhs(i, 1 + j);
      for (int i = 0; i < nc; i++) Z(i, j) = rhs(nu + i, 1 + j);
    }
    for (int j = 0; j < nth; j++) {
      for (int i = 0; i < nu; i++) Kth(i, j) = rhs(i, 1 + nx + j);
      for (int i = 0; i < nc; i++) Zth(i, j) = rhs(nu + i, 1 + nx + j);
    }
  }

  vc.Pmat = model.Q;                               // :134-140
  gemm_tn(1.0, model.C, Z, 1.0, vc.Pmat);
  vc.pvec = model.q;
  gemm_tn(1.0, model.C, zff, 1.0, vc.pvec);
  if (nu > 0) {
    gemm_nn(1.0, model.S, K, 1.0, vc.Pmat);
    gemm_nn(1.0, model.S, kff, 1.0, vc.pvec);
  }
  if (nth > 0) {                                   // :142-149
    vc.Vxt = model.Gx;
    gemm_tn(1.0, K, model.Gu, 1.0, vc.Vxt);
    vc.Vtt = model.Gth;
    gemm_tn(1.0, model.Gu, Kth, 1.0, vc.Vtt);
    vc.vt = model.gamma;
    gemm_tn(1.0, model.Gu, kff, 1.0, vc.vt);
  }
  d.ff.setZero(); d.fb.setZero(); d.fth.setZero();
  put_rows(d.ff, 0, kff);  put_rows(d.ff, nu, zff);
  put_rows(d.fb, 0, K);    put_rows(d.fb, nu, Z);
  put_rows(d.fth, 0, Kth); put_rows(d.fth, nu, Zth);
}

// riccati-kernel.hxx:167-300
void stageKernelSolve(const Knot &model, StageFactor &d, Value &vn, double mudyn, double mueq) {
  // workspaces are thread-local and reused: nothing is allocated per knot after the first one
  // (ALIGATOR_NOMALLOC_SCOPED, proximal-riccati.hxx:35)
  const int nx = model.nx, nu = model.nu, nc = model.nc, nth = model.nth;
  // step 1: E^{-1} by partial-pivot LU applied to the identity         (:172-174)
  d.Efact.compute(model.E);
  d.Einv.resize(nx, nx);
  for (int i = 0; i < nx; i++) d.Einv(i, i) = 1.0;
  d.Efact.solveInPlace(d.Einv);
  static thread_local Mat ptilde; ptilde.resize(nx);                                                      // (:175-177)
  gemm_tn(-1.0, d.Einv, vn.pvec, 0.0, ptilde);
  static thread_local Mat EinvP; EinvP.resize(nx, nx);                                                   // (:178-180)
  gemm_tn(1.0, d.Einv, vn.Pmat, 0.0, EinvP);
  gemm_nn(1.0, EinvP, d.Einv, 0.0, d.Ptilde);
  symmetrize_from_lower(d.Ptilde);

  static thread_local Mat schur; schur.resize(nx, nx);                                                   // (:182-185)
  for (int j = 0; j < nx; j++)
    for (int i = 0; i < nx; i++) schur(i, j) = (i == j ? 1.0 : 0.0) + mudyn * d.Ptilde(i, j);
  d.schurChol.compute(schur);
  vn.Vxx = d.Ptilde;                                                   // (:186-190)
  vn.vx = ptilde;
  gemm_nn(1.0, d.Ptilde, model.f, 1.0, vn.vx);
  d.schurChol.solveInPlace(vn.vx);
  d.schurChol.solveInPlace(vn.Vxx);
  symmetrize_from_lower(vn.Vxx);

  gemm_tn(1.0, model.A, vn.Vxx, 0.0, d.AtV);                           // (:192-199)
  gemm_tn(1.0, model.B, vn.Vxx, 0.0, d.BtV);
  d.Qhat = model.Q;  gemm_nn(1.0, d.AtV, model.A, 1.0, d.Qhat);
  d.Rhat = model.R;  gemm_nn(1.0, d.BtV, model.B, 1.0, d.Rhat);
  d.Shat = model.S;  gemm_nn(1.0, d.AtV, model.B, 1.0, d.Shat);
  d.qhat = model.q;  gemm_tn(1.0, model.A, vn.vx, 1.0, d.qhat);
  d.rhat = model.r;  gemm_tn(1.0, model.B, vn.vx, 1.0, d.rhat);

  // reduced KKT [Rhat D^T; D -mueq I], lower triangle -> BK                (:202-208)
  static thread_local Mat kkt; kkt = assemble_reduced_kkt(d.Rhat, model.D, mueq);
  d.kktChol.compute(kkt);

  static thread_local Mat rhs; rhs.resize(nu + nc, 1 + nx);                                            // (:216-228)
  for (int i = 0; i < nu; i++) rhs(i, 0) = -d.rhat[i];
  for (int i = 0; i < nc; i++) rhs(nu + i, 0) = -model.d[i];
  for (int j = 0; j < nx; j++) {
    for (int i = 0; i < nu; i++) rhs(i, 1 + j) = -d.Shat(j, i);
    for (int i = 0; i < nc; i++) rhs(nu + i, 1 + j) = -model.C(i, j);
  }
  d.kktChol.solveInPlace(rhs);
  static thread_local Mat kff, zff, K, Z;
  kff.resize(nu); zff.resize(nc); K.resize(nu, nx); Z.resize(nc, nx);
  for (int i = 0; i < nu; i++) kff[i] = rhs(i, 0);
  for (int i = 0; i < nc; i++) zff[i] = rhs(nu + i, 0);
  for (int j = 0; j < nx; j++) {
    for (int i = 0; i < nu; i++) K(i, j) = rhs(i, 1 + j);
    for (int i = 0; i < nc; i++) Z(i, j) = rhs(nu + i, 1 + j);
  }

  // closed-loop dynamics                                               (:231-245)
  static thread_local Mat lff; lff = vn.vx;
  gemm_tn(1.0, d.BtV, kff, 1.0, lff);
  static thread_local Mat yff_pre; yff_pre = model.f;
  gemm_nn(1.0, model.B, kff, 1.0, yff_pre);
  axpy(-mudyn, lff, yff_pre);
  static thread_local Mat yff; yff.resize(nx);
  gemm_nn(-1.0, d.Einv, yff_pre, 0.0, yff);

  static thread_local Mat L; L.resize(nx, nx);
  gemm_nn(1.0, vn.Vxx, model.A, 0.0, L);
  gemm_tn(1.0, d.BtV, K, 1.0, L);
  static thread_local Mat A_pre; A_pre = model.A;
  gemm_nn(1.0, model.B, K, 1.0, A_pre);
  axpy(-mudyn, L, A_pre);
  static thread_local Mat Afb; Afb.resize(nx, nx);
  gemm_nn(-1.0, d.Einv, A_pre, 0.0, Afb);

  Value &vc = d.vm;                                                    // (:247-250)
  vc.Pmat = d.Qhat;
  gemm_nn(1.0, d.Shat, K, 1.0, vc.Pmat);
  gemm_tn(1.0, model.C, Z, 1.0, vc.Pmat);
  vc.pvec = d.qhat;
  gemm_nn(1.0, d.Shat, kff, 1.0, vc.pvec);
  gemm_tn(1.0, model.C, zff, 1.0, vc.pvec);

  static thread_local Mat Kth, Zth, Lth, Yth;
  Kth.resize(nu, nth); Zth.resize(nc, nth); Lth.resize(nx, nth); Yth.resize(nx, nth);
  if (nth > 0) {                                                       // (:256-299)
    gemm_tn(-1.0, d.Einv, vn.Vxt, 0.0, Lth);      // Pxttilde = -E^{-T} Vxt
    d.schurChol.solveInPlace(Lth);                // Lambda^{-1} Pxttilde
    d.Guhat = model.Gu;
    gemm_tn(1.0, model.B, Lth, 1.0, d.Guhat);
    static thread_local Mat rth; rth.resize(nu + nc, nth);
    for (int j = 0; j < nth; j++)
      for (int i = 0; i < nu; i++) rth(i, j) = -d.Guhat(i, j);   // Zth rhs = 0 (:272)
    d.kktChol.solveInPlace(rth);
    Kth = get_rows(rth, 0, nu);
    Zth = get_rows(rth, nu, nc);
    gemm_tn(1.0, d.BtV, Kth, 1.0, Lth);
    static thread_local Mat Yth_pre; Yth_pre.resize(nx, nth);
    gemm_nn(1.0, model.B, Kth, 0.0, Yth_pre);
    axpy(-mudyn, Lth, Yth_pre);
    gemm_nn(-1.0, d.Einv, Yth_pre, 0.0, Yth);

    vc.vt = vn.vt;                                                     // (:286-289)
    axpy(1.0, model.gamma, vc.vt);
    gemm_tn(1.0, model.Gu, kff, 1.0, vc.vt);
    gemm_tn(1.0, vn.Vxt, yff, 1.0, vc.vt);
    vc.Vxt = model.Gx;                                                 // (:292-294)
    gemm_tn(1.0, K, model.Gu, 1.0, vc.Vxt);
    gemm_tn(1.0, Afb, vn.Vxt, 1.0, vc.Vxt);
    vc.Vtt = model.Gth;                                                // (:296-298)
    axpy(1.0, vn.Vtt, vc.Vtt);
    gemm_tn(1.0, model.Gu, Kth, 1.0, vc.Vtt);
    gemm_tn(1.0, vn.Vxt, Yth, 1.0, vc.Vtt);
  }

  put_rows(d.ff, 0, kff); put_rows(d.ff, nu, zff); put_rows(d.ff, nu + nc, lff);
  put_rows(d.ff, nu + nc + nx, yff);
  put_rows(d.fb, 0, K); put_rows(d.fb, nu, Z); put_rows(d.fb, nu + nc, L);
  put_rows(d.fb, nu + nc + nx, Afb);
  put_rows(d.fth, 0, Kth); put_rows(d.fth, nu, Zth); put_rows(d.fth, nu + nc, Lth);
  put_rows(d.fth, nu + nc + nx, Yth);
}

// riccati-kernel.hxx:64-88
bool backwardImpl(const std::vector<Knot> &stages, int beg, int end, double mudyn, double mueq,
                  std::vector<StageFactor> &datas) {
  if (end <= beg) return true;
  const int N = end - 1;
  terminalSolve(stages[N], mueq, datas[N]);
  for (int t = N - 1; t >= beg; t--)
    stageKernelSolve(stages[t], datas[t], datas[t + 1].vm, mudyn, mueq);
  return true;
}

// riccati-kernel.hxx:303-361
bool forwardImpl(const std::vector<Knot> &stages, const std::vector<StageFactor> &datas, int beg,
                 int end, std::vector<Mat> &xs, std::vector<Mat> &us, std::vector<Mat> &vs,
                 std::vector<Mat> &lbdas, const Mat *theta) {
  const int N = end - 1;
  for (int t = beg; t <= N; t++) {
    const StageFactor &d = datas[t];
    const Knot &model = stages[t];
    const int nx = model.nx, nu = model.nu, nc = model.nc, nth = model.nth;
    const bool par = nth > 0 && theta != nullptr;
    auto row_apply = [&](int r0, int nr, Mat &out) {
      for (int i = 0; i < nr; i++) {
        double s = d.ff[r0 + i];
        for (int j = 0; j < nx; j++) s += d.fb(r0 + i, j) * xs[t][j];
        if (par)
          for (int j = 0; j < nth; j++) s += d.fth(r0 + i, j) * (*theta)[j];
        out[i] = s;
      }
    };
    if (nu > 0) row_apply(0, nu, us[t]);
    row_apply(nu, nc, vs[t]);
    if (t == N) break;
    row_apply(nu + nc, nx, lbdas[t + 1]);
    row_apply(nu + nc + nx, nx, xs[t + 1]);
  }
  return true;
}

// proximal-riccati.hxx:12-30
ProximalRiccatiSolver::ProximalRiccatiSolver(const Problem &p) : problem(&p) {
  const int N = p.horizon();
  datas.reserve(N + 1);
  for (int t = 0; t <= N; t++) {
    const Knot &k = p.stages[t];
    datas.emplace_back(k.nx, k.nu, k.nc, k.nth);
  }
  const int nth = p.ntheta();
  thGrad.resize(nth);
  thHess.resize(nth, nth);
}

// proximal-riccati.hxx:32-64
bool ProximalRiccatiSolver::backward(double mudyn, double mueq) {
  const Problem &p = *problem;
  bool ret = backwardImpl(p.stages, 0, (int)p.stages.size(), mudyn, mueq, datas);
  Value &vinit = datas[0].vm;
  vinit.Vxx = vinit.Pmat;
  vinit.vx = vinit.pvec;
  const int nx = p.stages[0].nx, nc0 = p.nc0(), nth = p.ntheta();
  Mat kkt(nx + nc0, nx + nc0);
  for (int j = 0; j < nx; j++)
    for (int i = 0; i < nx; i++) kkt(i, j) = vinit.Vxx(i, j);
  for (int j = 0; j < nx; j++)
    for (int i = 0; i < nc0; i++) { kkt(nx + i, j) = p.G0(i, j); kkt(j, nx + i) = p.G0(i, j); }
  for (int i = 0; i < nc0; i++) kkt(nx + i, nx + i) = -mudyn;
  kkt0.chol.compute(kkt);
  kkt0.ff.resize(nx + nc0);
  for (int i = 0; i < nx; i++) kkt0.ff[i] = -vinit.vx[i];
  for (int i = 0; i < nc0; i++) kkt0.ff[nx + i] = -p.g0[i];
  kkt0.chol.solveInPlace(kkt0.ff);
  kkt0.fth.resize(nx + nc0, nth);
  for (int j = 0; j < nth; j++)
    for (int i = 0; i < nx; i++) kkt0.fth(i, j) = -vinit.Vxt(i, j);
  if (nth > 0) kkt0.chol.solveInPlace(kkt0.fth);
  // thGrad = vt + Vxt^T x0ff ; thHess = Vtt + Vxt^T fth_x        (:59-61)
  thGrad = vinit.vt;
  thHess = vinit.Vtt;
  for (int a = 0; a < nth; a++) {
    double s = 0.0;
    for (int i = 0; i < nx; i++) s += vinit.Vxt(i, a) * kkt0.ff[i];
    thGrad[a] += s;
    for (int b = 0; b < nth; b++) {
      double h = 0.0;
      for (int i = 0; i < nx; i++) h += vinit.Vxt(i, a) * kkt0.fth(i, b);
      thHess(a, b) += h;
    }
  }
  return ret;
}

// proximal-riccati.hxx:66-78 + riccati-kernel.hxx:153-164
bool ProximalRiccatiSolver::forward(std::vector<Mat> &xs, std::vector<Mat> &us,
                                    std::vector<Mat> &vs, std::vector<Mat> &lbdas,
                                    const Mat *theta) const {
  const Problem &p = *problem;
  const int nx = p.stages[0].nx, nc0 = p.nc0(), nth = p.ntheta();
  for (int i = 0; i < nx; i++) xs[0][i] = kkt0.ff[i];
  for (int i = 0; i < nc0; i++) lbdas[0][i] = kkt0.ff[nx + i];
  if (theta) {
    for (int j = 0; j < nth; j++) {
      for (int i = 0; i < nx; i++) xs[0][i] += kkt0.fth(i, j) * (*theta)[j];
      for (int i = 0; i < nc0; i++) lbdas[0][i] += kkt0.fth(nx + i, j) * (*theta)[j];
    }
  }
  return forwardImpl(p.stages, datas, 0, (int)p.stages.size(), xs, us, vs, lbdas, theta);
}

} // namespace gar_oracle
