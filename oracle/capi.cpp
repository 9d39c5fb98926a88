// oracle/capi.cpp — TEST INFRASTRUCTURE ONLY.
// Flat extern "C" entry points (ctypes-friendly) over the oracle.  Problems, gains and
// value functions travel in the packed record format of include/garb200_layout.h, i.e.
// the oracle and the CUDA path are fed THE SAME BYTES.
#include "gar_oracle.hpp"
#include "../include/garb200_layout.h"
#include <chrono>
#include <cstdio>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace gar_oracle;

namespace {

Knot unpack_knot(const garb200_knot_layout &L, int nu, int nc, const double *rec) {
  Knot k(L.nx, nu, nc, L.nth);
  k.Q.load(rec + L.Q); k.S.load(rec + L.S); k.R.load(rec + L.R);
  k.q.load(rec + L.q); k.r.load(rec + L.r);
  k.A.load(rec + L.A); k.B.load(rec + L.B); k.E.load(rec + L.E); k.f.load(rec + L.f);
  k.C.load(rec + L.C); k.D.load(rec + L.D); k.d.load(rec + L.d);
  k.Gth.load(rec + L.Gth); k.Gx.load(rec + L.Gx); k.Gu.load(rec + L.Gu);
  k.gamma.load(rec + L.gamma);
  return k;
}

Problem unpack_problem(int nx, int nu, int nc, int nth, int T, int nc0, const int *nu_t,
                       const int *nc_t, const double *knots, const double *G0,
                       const double *g0) {
  const garb200_knot_layout L = garb200_make_knot_layout(nx, nu, nc, nth);
  Problem p;
  p.stages.reserve(T + 1);
  for (int t = 0; t <= T; t++)
    p.stages.push_back(unpack_knot(L, nu_t[t], nc_t[t], knots + (size_t)t * L.stride));
  p.G0.resize(nc0, nx); p.G0.load(G0);
  p.g0.resize(nc0);     p.g0.load(g0);
  return p;
}

void store_rowmajor(const Mat &M, double *dst) {
  for (int i = 0; i < M.r; i++)
    for (int j = 0; j < M.c; j++) dst[(size_t)i * M.c + j] = M(i, j);
}

void export_factors(const std::vector<StageFactor> &datas, int nx, int nu, int nc, int nth_layout,
                    double *gains, double *values) {
  const garb200_gain_layout G = garb200_make_gain_layout(nx, nu, nc, nth_layout);
  const garb200_value_layout V = garb200_make_value_layout(nx, nth_layout);
  for (size_t t = 0; t < datas.size(); t++) {
    const StageFactor &d = datas[t];
    if (gains) {
      double *g = gains + t * G.stride;
      std::fill(g, g + G.stride, 0.0);
      d.ff.store(g + G.ff);
      store_rowmajor(d.fb, g + G.fb);
      if (d.nth > 0) store_rowmajor(d.fth, g + G.fth);
    }
    if (values) {
      double *v = values + t * V.stride;
      std::fill(v, v + V.stride, 0.0);
      d.vm.Pmat.store(v + V.Pmat); d.vm.pvec.store(v + V.pvec);
      d.vm.Vxx.store(v + V.Vxx);   d.vm.vx.store(v + V.vx);
      if (d.nth > 0) { d.vm.Vxt.store(v + V.Vxt); d.vm.Vtt.store(v + V.Vtt); d.vm.vt.store(v + V.vt); }
    }
  }
}

// trajectory layout: xs (T+1)*nx | us (T+1)*nu | vs (T+1)*nc | lbdas nc0 + T*nx
void export_traj(const std::vector<Mat> &xs, const std::vector<Mat> &us, const std::vector<Mat> &vs,
                 const std::vector<Mat> &lbdas, int nx, int nu, int nc, int nc0, int T,
                 double *oxs, double *ous, double *ovs, double *olb) {
  for (int t = 0; t <= T; t++) {
    for (int i = 0; i < nx; i++) oxs[(size_t)t * nx + i] = xs[t][i];
    for (int i = 0; i < nu; i++) ous[(size_t)t * nu + i] = 0.0;
    if (t < (int)us.size())
      for (int i = 0; i < us[t].r; i++) ous[(size_t)t * nu + i] = us[t][i];
    for (int i = 0; i < nc; i++) ovs[(size_t)t * nc + i] = 0.0;
    for (int i = 0; i < vs[t].r; i++) ovs[(size_t)t * nc + i] = vs[t][i];
  }
  for (int i = 0; i < nc0; i++) olb[i] = lbdas[0][i];
  for (int t = 1; t <= T; t++)
    for (int i = 0; i < nx; i++) olb[nc0 + (size_t)(t - 1) * nx + i] = lbdas[t][i];
}

void import_traj(const Problem &p, const double *ixs, const double *ius, const double *ivs,
                 const double *ilb, int nx, int nu, int nc, int nc0, std::vector<Mat> &xs,
                 std::vector<Mat> &us, std::vector<Mat> &vs, std::vector<Mat> &lbdas) {
  const int T = p.horizon();
  lqrInitializeSolution(p, xs, us, vs, lbdas);
  for (int t = 0; t <= T; t++) {
    for (int i = 0; i < nx; i++) xs[t][i] = ixs[(size_t)t * nx + i];
    if (t < (int)us.size())
      for (int i = 0; i < us[t].r; i++) us[t][i] = ius[(size_t)t * nu + i];
    for (int i = 0; i < vs[t].r; i++) vs[t][i] = ivs[(size_t)t * nc + i];
  }
  if (us.size() < (size_t)T + 1) us.resize(T + 1);  // keep us[T] addressable (size 0)
  for (int i = 0; i < nc0; i++) lbdas[0][i] = ilb[i];
  for (int t = 1; t <= T; t++)
    for (int i = 0; i < nx; i++) lbdas[t][i] = ilb[nc0 + (size_t)(t - 1) * nx + i];
}

} // namespace

extern "C" {

/// Layout queries (so Python never re-derives include/garb200_layout.h by hand).
/// knot: [Q,S,R,q,r,A,B,E,f,C,D,d,Gth,Gx,Gu,gamma,stride]; gain: [nrows,ff,fb,fth,stride];
/// value: [Pmat,pvec,Vxx,vx,Vxt,Vtt,vt,stride]
void gar_oracle_layouts(int nx, int nu, int nc, int nth, int *knot, int *gain, int *value) {
  const garb200_knot_layout L = garb200_make_knot_layout(nx, nu, nc, nth);
  const int k[17] = {L.Q, L.S, L.R, L.q, L.r, L.A, L.B, L.E, L.f, L.C, L.D, L.d,
                     L.Gth, L.Gx, L.Gu, L.gamma, L.stride};
  for (int i = 0; i < 17; i++) knot[i] = k[i];
  const garb200_gain_layout G = garb200_make_gain_layout(nx, nu, nc, nth);
  gain[0] = G.nrows; gain[1] = G.ff; gain[2] = G.fb; gain[3] = G.fth; gain[4] = G.stride;
  const garb200_value_layout V = garb200_make_value_layout(nx, nth);
  const int v[8] = {V.Pmat, V.pvec, V.Vxx, V.vx, V.Vxt, V.Vtt, V.vt, V.stride};
  for (int i = 0; i < 8; i++) value[i] = v[i];
}

/// ProximalRiccatiSolver backward + forward on one packed problem.
/// Any output pointer may be NULL.  kkt0_fth / fth blocks are row-major.
int gar_oracle_prox_solve(int nx, int nu, int nc, int nth, int T, int nc0, const int *nu_t,
                          const int *nc_t, const double *knots, const double *G0,
                          const double *g0, double mudyn, double mueq, const double *theta,
                          double *gains, double *values, double *kkt0_ff, double *kkt0_fth,
                          double *thGrad, double *thHess, double *oxs, double *ous,
                          double *ovs, double *olb, double *kkt_err) {
  Problem p = unpack_problem(nx, nu, nc, nth, T, nc0, nu_t, nc_t, knots, G0, g0);
  ProximalRiccatiSolver solver(p);
  solver.backward(mudyn, mueq);
  std::vector<Mat> xs, us, vs, lbdas;
  lqrInitializeSolution(p, xs, us, vs, lbdas);
  if (us.size() < (size_t)T + 1) us.resize(T + 1);
  Mat th(nth);
  if (theta) th.load(theta);
  solver.forward(xs, us, vs, lbdas, theta ? &th : nullptr);
  export_factors(solver.datas, nx, nu, nc, nth, gains, values);
  if (kkt0_ff) solver.kkt0.ff.store(kkt0_ff);
  if (kkt0_fth && nth > 0) store_rowmajor(solver.kkt0.fth, kkt0_fth);
  if (thGrad && nth > 0) solver.thGrad.store(thGrad);
  if (thHess && nth > 0) solver.thHess.store(thHess);
  if (oxs) export_traj(xs, us, vs, lbdas, nx, nu, nc, nc0, T, oxs, ous, ovs, olb);
  if (kkt_err) {
    auto e = lqrComputeKktError(p, xs, us, vs, lbdas, mudyn, mueq, theta ? &th : nullptr);
    kkt_err[0] = e[0]; kkt_err[1] = e[1]; kkt_err[2] = e[2];
  }
  return 0;
}

/// RiccatiSolverDense backward + forward on one packed problem.  gv: [T+1][nc * nth] column-major blocks
/// with the solver's MAX nc as leading dimension, or NULL (zeros) — Gv is not part of the packed knot record.
int gar_oracle_dense_solve(int nx, int nu, int nc, int nth, int T, int nc0, const int *nu_t,
                           const int *nc_t, const double *knots, const double *G0, const double *g0,
                           const double *gv, double mudyn, double mueq, const double *theta, double *gains,
                           double *values, double *kkt0_ff, double *kkt0_fth, double *thGrad, double *thHess,
                           double *oxs, double *ous, double *ovs, double *olb, double *kkt_err) {
  Problem p = unpack_problem(nx, nu, nc, nth, T, nc0, nu_t, nc_t, knots, G0, g0);
  if (gv)
    for (int t = 0; t <= T; t++) {
      Knot &k = p.stages[t];
      for (int j = 0; j < nth; j++)
        for (int i = 0; i < k.nc; i++) k.Gv(i, j) = gv[(size_t)t * nc * nth + i + (size_t)j * nc];
    }
  RiccatiSolverDense solver(p);
  solver.backward(mudyn, mueq);
  std::vector<Mat> xs, us, vs, lbdas;
  lqrInitializeSolution(p, xs, us, vs, lbdas);
  if (us.size() < (size_t)T + 1) us.resize(T + 1);
  Mat th(nth);
  if (theta) th.load(theta);
  solver.forward(xs, us, vs, lbdas, theta ? &th : nullptr);
  export_factors(solver.datas, nx, nu, nc, nth, gains, values);
  if (kkt0_ff) solver.kkt0.ff.store(kkt0_ff);
  if (kkt0_fth && nth > 0) store_rowmajor(solver.kkt0.fth, kkt0_fth);
  if (thGrad && nth > 0) solver.thGrad.store(thGrad);
  if (thHess && nth > 0) solver.thHess.store(thHess);
  if (oxs) export_traj(xs, us, vs, lbdas, nx, nu, nc, nc0, T, oxs, ous, ovs, olb);
  if (kkt_err) {   // (lqrComputeKktError has no Gv term, utils.hxx:94-192: meaningful for Gv = 0 only)
    auto e = lqrComputeKktError(p, xs, us, vs, lbdas, mudyn, mueq, theta ? &th : nullptr);
    kkt_err[0] = e[0]; kkt_err[1] = e[1]; kkt_err[2] = e[2];
  }
  return 0;
}

/// ParallelRiccatiSolver (num_legs >= 2) on one packed problem with nth = 0.
/// gains/values use the layout built with nth = nx (the legs' parameter dimension).
/// info[0] = refinement steps taken, info[1] = block-tridiagonal solve ok.
int gar_oracle_parallel_solve(int nx, int nu, int nc, int T, int nc0, const int *nu_t,
                              const int *nc_t, const double *knots, const double *G0,
                              const double *g0, double mudyn, double mueq, int num_legs,
                              int collapse, double *gains, double *values, double *oxs,
                              double *ous, double *ovs, double *olb, double *kkt_err,
                              int *info) {
  if (num_legs < 2) return 1;
  Problem p = unpack_problem(nx, nu, nc, 0, T, nc0, nu_t, nc_t, knots, G0, g0);
  const Problem pref = p; // un-mutated copy for the KKT residual
  ParallelRiccatiSolver solver(p, num_legs);
  solver.backward(mudyn, mueq);
  std::vector<Mat> xs, us, vs, lbdas;
  lqrInitializeSolution(pref, xs, us, vs, lbdas);
  if (us.size() < (size_t)T + 1) us.resize(T + 1);
  solver.forward(xs, us, vs, lbdas);
  if (collapse) solver.collapseFeedback();
  export_factors(solver.datas, nx, nu, nc, nx, gains, values);
  if (oxs) export_traj(xs, us, vs, lbdas, nx, nu, nc, nc0, T, oxs, ous, ovs, olb);
  if (kkt_err) {
    auto e = lqrComputeKktError(pref, xs, us, vs, lbdas, mudyn, mueq, nullptr);
    kkt_err[0] = e[0]; kkt_err[1] = e[1]; kkt_err[2] = e[2];
  }
  if (info) { info[0] = solver.refinementStepsTaken; info[1] = solver.tridiagOk ? 1 : 0; }
  return 0;
}

/// lqrComputeKktError on a packed problem + packed trajectory.
int gar_oracle_kkt_error(int nx, int nu, int nc, int nth, int T, int nc0, const int *nu_t,
                         const int *nc_t, const double *knots, const double *G0,
                         const double *g0, double mudyn, double mueq, const double *theta,
                         const double *ixs, const double *ius, const double *ivs,
                         const double *ilb, double *kkt_err) {
  Problem p = unpack_problem(nx, nu, nc, nth, T, nc0, nu_t, nc_t, knots, G0, g0);
  std::vector<Mat> xs, us, vs, lbdas;
  import_traj(p, ixs, ius, ivs, ilb, nx, nu, nc, nc0, xs, us, vs, lbdas);
  Mat th(nth);
  if (theta) th.load(theta);
  auto e = lqrComputeKktError(p, xs, us, vs, lbdas, mudyn, mueq, theta ? &th : nullptr);
  kkt_err[0] = e[0]; kkt_err[1] = e[1]; kkt_err[2] = e[2];
  return 0;
}

int gar_oracle_num_rows(int nx, int T, int nc0, const int *nu_t, const int *nc_t) {
  int n = nc0;
  for (int t = 0; t <= T; t++) { n += nx + nu_t[t] + nc_t[t]; if (t != T) n += nx; }
  return n;
}

/// lqrDenseMatrix: mat (nrows x nrows, column-major == row-major since symmetric), rhs.
int gar_oracle_dense_kkt(int nx, int nu, int nc, int T, int nc0, const int *nu_t,
                         const int *nc_t, const double *knots, const double *G0,
                         const double *g0, double mudyn, double mueq, double *mat,
                         double *rhs) {
  Problem p = unpack_problem(nx, nu, nc, 0, T, nc0, nu_t, nc_t, knots, G0, g0);
  Mat M, r;
  if (!lqrDenseMatrix(p, mudyn, mueq, M, r)) return 1;
  M.store(mat);
  r.store(rhs);
  return 0;
}

/// Factor/solve helpers exposed for unit tests of the restated factorizations.
int gar_oracle_bk_solve(int n, int nrhs, const double *A, double *X, int *pivots) {
  Mat a(n, n), x(n, nrhs);
  a.load(A); x.load(X);
  BunchKaufman bk;
  bk.compute(a);
  if (!bk.ok) return 1;
  bk.solveInPlace(x);
  x.store(X);
  if (pivots) for (int i = 0; i < n; i++) pivots[i] = bk.piv[i];
  return 0;
}
int gar_oracle_lu_solve(int n, int nrhs, const double *A, double *X) {
  Mat a(n, n), x(n, nrhs);
  a.load(A); x.load(X);
  PartialPivLU lu;
  lu.compute(a);
  if (!lu.ok) return 1;
  lu.solveInPlace(x);
  x.store(X);
  return 0;
}
int gar_oracle_llt_solve(int n, int nrhs, const double *A, double *X) {
  Mat a(n, n), x(n, nrhs);
  a.load(A); x.load(X);
  LLT llt;
  llt.compute(a);
  if (!llt.ok) return 1;
  llt.solveInPlace(x);
  x.store(X);
  return 0;
}

/// Symmetric block-tridiagonal solve on equally sized blocks (block-matrix.cpp:54-137 test).
/// diag: nb blocks (bs x bs), super: nb-1 blocks, rhs: nb*bs; column-major blocks.
int gar_oracle_blocktridiag_solve(int nb, int bs, const double *diag_, const double *super_,
                                  double *rhs_) {
  std::vector<Mat> diag(nb, Mat(bs, bs)), super(nb - 1, Mat(bs, bs)), sub(nb - 1), rhs(nb, Mat(bs));
  for (int i = 0; i < nb; i++) { diag[i].load(diag_ + (size_t)i * bs * bs); rhs[i].load(rhs_ + (size_t)i * bs); }
  for (int i = 0; i + 1 < nb; i++) { super[i].load(super_ + (size_t)i * bs * bs); sub[i] = transpose(super[i]); }
  std::vector<BunchKaufman> facs(nb);
  if (!symmetricBlockTridiagSolve(sub, diag, super, rhs, facs)) return 1;
  for (int i = 0; i < nb; i++) rhs[i].store(rhs_ + (size_t)i * bs);
  return 0;
}

/// Timed CPU baseline: `nbatch` independent packed problems, `reps` x (backward + forward)
/// each, OpenMP over problems with `nthreads` threads (0 = all).  Mirrors
/// /root/reference/bench/gar-riccati.cpp:18-32: problem + solver built outside the timed
/// loop.  num_legs >= 2 times the ParallelRiccatiSolver instead (OpenMP over legs).
/// Returns seconds for the whole timed region; *threads_used reports the team size.
double gar_oracle_bench(int nx, int nu, int nc, int nth, int T, int nc0, const int *nu_t,
                        const int *nc_t, int nbatch, const double *knots, const double *G0,
                        const double *g0, double mudyn, double mueq, const double *theta,
                        int num_legs, int reps, int nthreads, int *threads_used) {
  const garb200_knot_layout L = garb200_make_knot_layout(nx, nu, nc, nth);
  const size_t pstride = (size_t)(T + 1) * L.stride;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
  nthreads = 1;
#endif
  if (threads_used) *threads_used = nthreads;
  std::vector<Problem> probs;
  probs.reserve(nbatch);
  for (int b = 0; b < nbatch; b++)
    probs.push_back(unpack_problem(nx, nu, nc, nth, T, nc0, nu_t, nc_t, knots + b * pstride,
                                   G0 + (size_t)b * nc0 * nx, g0 + (size_t)b * nc0));
  Mat th(nth);
  if (theta) th.load(theta);
  double secs = 0.0;
  if (num_legs >= 2) {
    // parallel-in-time: one problem at a time, legs in parallel
    std::vector<ParallelRiccatiSolver> solvers;
    solvers.reserve(nbatch);
    for (int b = 0; b < nbatch; b++) solvers.emplace_back(probs[b], num_legs);
    std::vector<Mat> xs, us, vs, lbdas;
#ifdef _OPENMP
    omp_set_num_threads(nthreads);
#endif
    auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < reps; r++)
      for (int b = 0; b < nbatch; b++) {
        if (r == 0) { lqrInitializeSolution(probs[b], xs, us, vs, lbdas); if (us.size() < (size_t)T + 1) us.resize(T + 1); }
        solvers[b].backward(mudyn, mueq);
        solvers[b].forward(xs, us, vs, lbdas);
      }
    secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  } else {
    std::vector<ProximalRiccatiSolver> solvers;
    solvers.reserve(nbatch);
    for (int b = 0; b < nbatch; b++) solvers.emplace_back(probs[b]);
    std::vector<std::array<std::vector<Mat>, 4>> sols(nbatch);
    for (int b = 0; b < nbatch; b++) {
      lqrInitializeSolution(probs[b], sols[b][0], sols[b][1], sols[b][2], sols[b][3]);
      if (sols[b][1].size() < (size_t)T + 1) sols[b][1].resize(T + 1);
    }
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
    for (int b = 0; b < nbatch; b++)
      for (int r = 0; r < reps; r++) {
        solvers[b].backward(mudyn, mueq);
        solvers[b].forward(sols[b][0], sols[b][1], sols[b][2], sols[b][3], theta ? &th : nullptr);
      }
    secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }
  return secs;
}

} // extern "C"
