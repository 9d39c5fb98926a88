// oracle/parallel.cpp — TEST INFRASTRUCTURE ONLY.
// Restatement of ParallelRiccatiSolver (parallel-in-time leg condensation) and of the
// symmetric block-tridiagonal solver it uses.
//   /root/reference/include/aligator/gar/parallel-solver.{hpp,hxx}
//   /root/reference/include/aligator/gar/block-tridiagonal.hpp
#include "gar_oracle.hpp"
#include <stdexcept>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace gar_oracle {

static Mat neg_copy(const Mat &X) { Mat Y = X; scale(Y, -1.0); return Y; }

// block-tridiagonal.hpp:50-75 :  c <- beta*c + A*b
void blockTridiagMatMul(const std::vector<Mat> &sub, const std::vector<Mat> &diag,
                        const std::vector<Mat> &super, const std::vector<Mat> &b,
                        std::vector<Mat> &c, double beta) {
  const size_t N = super.size();
  for (auto &ci : c) scale(ci, beta);
  gemm_nn(1.0, diag[0], b[0], 1.0, c[0]);
  gemm_nn(1.0, super[0], b[1], 1.0, c[0]);
  for (size_t i = 1; i < N; i++) {
    gemm_nn(1.0, sub[i - 1], b[i - 1], 1.0, c[i]);
    gemm_nn(1.0, diag[i], b[i], 1.0, c[i]);
    gemm_nn(1.0, super[i], b[i + 1], 1.0, c[i]);
  }
  gemm_nn(1.0, sub[N - 1], b[N - 1], 1.0, c[N]);
  gemm_nn(1.0, diag[N], b[N], 1.0, c[N]);
}

// block-tridiagonal.hpp:80-137 (up-looking block LDL^T, in place)
bool symmetricBlockTridiagSolve(std::vector<Mat> &sub, std::vector<Mat> &diag,
                                const std::vector<Mat> &super, std::vector<Mat> &rhs,
                                std::vector<BunchKaufman> &facs) {
  if (sub.size() != super.size() || diag.size() != super.size() + 1 ||
      rhs.size() != diag.size())
    return false;
  const size_t N = super.size();
  for (size_t i = N; i-- > 0;) {
    BunchKaufman &ldl = facs[i + 1];
    ldl.compute(diag[i + 1]);
    if (!ldl.ok) return false;
    ldl.solveInPlace(rhs[i + 1]);
    const Mat &B = super[i];
    Mat &Cm = sub[i];
    gemm_nn(-1.0, B, rhs[i + 1], 1.0, rhs[i]);
    ldl.solveInPlace(Cm);                     // U^T = D_{i+1}^{-1} B_{i+1}^T
    gemm_nn(-1.0, B, Cm, 1.0, diag[i]);
  }
  BunchKaufman &ldl0 = facs[0];
  ldl0.compute(diag[0]);
  if (!ldl0.ok) return false;
  ldl0.solveInPlace(rhs[0]);
  for (size_t i = 0; i < N; i++) gemm_nn(-1.0, sub[i], rhs[i], 1.0, rhs[i + 1]);
  return true;
}

// block-tridiagonal.hpp:144-179
void blockTridiagRefinementStep(const std::vector<Mat> &upFacs, const std::vector<Mat> &super,
                                const std::vector<BunchKaufman> &facs, std::vector<Mat> &rhs) {
  const size_t N = super.size();
  for (size_t i = N; i-- > 0;) {
    facs[i + 1].solveInPlace(rhs[i + 1]);
    gemm_nn(-1.0, super[i], rhs[i + 1], 1.0, rhs[i]);
  }
  facs[0].solveInPlace(rhs[0]);
  for (size_t i = 0; i < N; i++) gemm_nn(-1.0, upFacs[i], rhs[i], 1.0, rhs[i + 1]);
}

// parallel-solver.hxx:18-61, :240-267
ParallelRiccatiSolver::ParallelRiccatiSolver(Problem &p, int num_threads)
    : numThreads(num_threads), problem(&p) {
  if (num_threads < 2)                               // :25-29 (ALIGATOR_RUNTIME_ERROR)
    throw std::runtime_error("num_threads should be greater than or equal to 2.");
  const int N = p.horizon();
  for (int i = 0; i < num_threads; i++) {
    auto [start, end] = get_work(N, i, num_threads);
    const bool last_leg = (i == num_threads - 1);
    for (int t = start; t < end; t++) {              // allocateLeg :51-61
      Knot &knot = p.stages[t];
      if (!last_leg) knot.addParameterization(knot.nx);
      datas.emplace_back(knot.nx, knot.nu, knot.nc, knot.nth);
    }
  }
  std::vector<int> dims{p.nc0(), p.stages.front().nx};
  for (int i = 0; i < num_threads - 1; i++) {
    auto [i0, i1] = get_work(N, i, num_threads);
    dims.push_back(p.stages[i0].nx);
    dims.push_back(p.stages[i1 - 1].nx);
  }
  const size_t nb = dims.size();
  diag.resize(nb); rhs.resize(nb); sol.resize(nb); err.resize(nb);
  diagFacs.resize(nb); upFacs.resize(nb ? nb - 1 : 0); ldlt.resize(nb);
  super.resize(nb - 1); sub.resize(nb - 1);
  for (size_t i = 0; i < nb; i++) {
    diag[i].resize(dims[i], dims[i]);
    rhs[i].resize(dims[i]); sol[i].resize(dims[i]); err[i].resize(dims[i]);
    if (i + 1 < nb) {
      super[i].resize(dims[i], dims[i + 1]);
      sub[i].resize(dims[i + 1], dims[i]);
    }
  }
}

// parallel-solver.hpp:43-51
void ParallelRiccatiSolver::setupKnot(Knot &knot, double mudyn) {
  knot.Gx = transpose(knot.A);
  knot.Gu = transpose(knot.B);
  knot.Gth.setZero();
  for (int i = 0; i < knot.Gth.r; i++) knot.Gth(i, i) = -mudyn;
  knot.gamma = knot.f;
}

// parallel-solver.hxx:63-109
void ParallelRiccatiSolver::assembleCondensedSystem(double mudyn) {
  const Problem &p = *problem;
  const auto &stages = p.stages;
  diag[0].setZero();
  for (int i = 0; i < diag[0].r; i++) diag[0](i, i) = -mudyn;
  super[0] = p.G0;
  diag[1] = datas[0].vm.Pmat;
  super[1] = datas[0].vm.Vxt;
  const int N = p.horizon();
  for (int i = 0; i < numThreads - 1; i++) {
    auto [i0, i1] = get_work(N, i, numThreads);
    const int ip1 = i + 1;
    diag[2 * ip1] = datas[i0].vm.Vtt;
    diag[2 * ip1 + 1] = datas[i1].vm.Pmat;
    super[2 * ip1] = stages[i1 - 1].E;
    if (ip1 + 1 < numThreads) super[2 * ip1 + 1] = datas[i1].vm.Vxt;
  }
  for (size_t i = 0; i < sub.size(); i++) sub[i] = transpose(super[i]);
  rhs[0] = neg_copy(p.g0);
  rhs[1] = neg_copy(datas[0].vm.pvec);
  for (int i = 0; i < numThreads - 1; i++) {
    auto [i0, i1] = get_work(N, i, numThreads);
    const int ip1 = i + 1;
    rhs[2 * ip1] = neg_copy(datas[i0].vm.vt);
    rhs[2 * ip1 + 1] = neg_copy(datas[i1].vm.pvec);
  }
}

// parallel-solver.hxx:111-179
bool ParallelRiccatiSolver::backward(double mudyn, double mueq) {
  Problem &p = *problem;
  const int N = p.horizon();
  for (int i = 0; i < numThreads - 1; i++) {
    const int end = get_work(N, i, numThreads)[1];
    setupKnot(p.stages[end - 1], mudyn);
  }
#pragma omp parallel for schedule(static, 1)
  for (int i = 0; i < numThreads; i++) {
    auto [beg, end] = get_work(N, i, numThreads);
    backwardImpl(p.stages, beg, end, mudyn, mueq, datas);
  }
  assembleCondensedSystem(mudyn);
  sol = rhs;
  diagFacs = diag;
  upFacs = sub;
  // NB the reference drops this bool (parallel-solver.hxx:148-151); it is kept here.
  tridiagOk = symmetricBlockTridiagSolve(sub, diag, super, sol, ldlt);
  diagFacs.swap(diag);
  upFacs.swap(sub);

  refinementStepsTaken = 0;
  constexpr int maxRefinementSteps = 5;
  for (int it = 0; it < maxRefinementSteps; it++) {
    // err <- rhs - A sol                                                (:159-163)
    // Reference quirk: on entry `condensedErr` holds whatever the previous call left
    // (uninitialised storage on the first call, the last residual after an early
    // return), not rhs; the next iteration repairs it.  The intended semantics
    // (err = rhs on entry) are restated; results agree to the refinement threshold.
    if (it == 0) err = rhs;
    blockTridiagMatMul(sub, diag, super, sol, err, -1.0);
    double resdl = 0.0;
    for (auto &e : err) { scale(e, -1.0); resdl = std::max(resdl, inf_norm(e)); }
    if (resdl <= condensedThreshold) return true;
    blockTridiagRefinementStep(upFacs, super, ldlt, err);
    for (size_t i = 0; i < sol.size(); i++) axpy(1.0, err[i], sol[i]);
    err = rhs;
    refinementStepsTaken++;
  }
  return true;
}

// parallel-solver.hxx:181-215
bool ParallelRiccatiSolver::forward(std::vector<Mat> &xs, std::vector<Mat> &us,
                                    std::vector<Mat> &vs, std::vector<Mat> &lbdas) const {
  const Problem &p = *problem;
  const int N = p.horizon();
  for (int i = 0; i < numThreads; i++) {
    const int i0 = get_work(N, i, numThreads)[0];
    lbdas[i0] = sol[2 * i];
    xs[i0] = sol[2 * i + 1];
  }
#pragma omp parallel for schedule(static, 1)
  for (int i = 0; i < numThreads; i++) {
    auto [beg, end] = get_work(N, i, numThreads);
    if (i < numThreads - 1) {
      const Mat theta = lbdas[end]; // set above from the condensed solution
      forwardImpl(p.stages, datas, beg, end, xs, us, vs, lbdas, &theta);
    } else {
      forwardImpl(p.stages, datas, beg, end, xs, us, vs, lbdas, nullptr);
    }
  }
  return true;
}

// parallel-solver.hpp:55-66 : K0 -= Kth0 * subdiagonal[1]
void ParallelRiccatiSolver::collapseFeedback() {
  StageFactor &d = datas[0];
  const Mat &Up1t = sub[1];
  for (int i = 0; i < d.nu; i++)
    for (int j = 0; j < d.nx; j++) {
      double s = 0.0;
      for (int l = 0; l < d.nth; l++) s += d.fth(i, l) * Up1t(l, j);
      d.fb(i, j) -= s;
    }
}

} // namespace gar_oracle
