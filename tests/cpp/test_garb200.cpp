// The reference's gar tests re-expressed on the C++ mirror (include/garb200.hpp):
//   short_horz_pb, one_knot_prob, random_long_problem   /root/reference/tests/gar/riccati.cpp:20-158
//   parallel_solver_class                               /root/reference/tests/gar/parallel.cpp:199-271
// Problem generator: restatement of tests/gar/test_util.cpp:13-66 (std::mt19937{42}).
// Exit code 0 = all assertions hold.  Needs a CUDA device (run by tests/test_cpp_mirror.py).
#include <cmath>
#include <cstdio>
#include <random>
#include "garb200.hpp"

using namespace garb200;
static std::mt19937 rng{42};
static std::normal_distribution<double> nrm;
static int fails = 0;
#define CHECK(cond, ...) do { if (!(cond)) { fails++; printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } } while (0)

static Mat wishart(int n, int p) {
  Mat W(n, p), out(n, n);
  for (auto &x : W.v) x = nrm(rng);
  for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) { double s = 0; for (int k = 0; k < p; k++) s += W(i, k) * W(j, k); out(i, j) = s; }
  return out;
}
static LqrKnot generate_knot(int nx, int nu, int nth, bool singular) {      // test_util.cpp:13-46
  LqrKnot k(nx, nu, 0, nx, nth);
  Mat qsr = wishart(nx + nu, nx + nu + 1);
  for (int i = 0; i < nx; i++) for (int j = 0; j < nx; j++) k.Q(i, j) = qsr(i, j);
  if (singular) for (int i = 0; i < nx / 2; i++) for (int j = 0; j < nx / 2; j++) k.Q(i, j) = 0.0;
  for (int i = 0; i < nx; i++) for (int j = 0; j < nu; j++) k.S(i, j) = qsr(i, nx + j);
  for (int i = 0; i < nu; i++) for (int j = 0; j < nu; j++) k.R(i, j) = qsr(nx + i, nx + j);
  for (auto &x : k.q) x = nrm(rng);
  for (auto &x : k.r) x = nrm(rng);
  for (auto &x : k.A.v) x = nrm(rng);
  std::uniform_real_distribution<double> uni(-1.0, 1.0);
  for (auto &x : k.B.v) x = uni(rng);
  for (auto &x : k.E.v) x = 1000.0 * nrm(rng);
  for (auto &x : k.f) x = 100.0 * nrm(rng);
  if (nth > 0) {
    for (auto &x : k.Gx.v) x = nrm(rng);
    for (auto &x : k.Gu.v) x = nrm(rng);
    k.Gth = wishart(nth, nth + 2);
    for (auto &x : k.gamma) x = nrm(rng);
  }
  return k;
}
static LqrProblem generate_problem(const Vec &x0, int horz, int nx, int nu, int nth = 0) {   // test_util.cpp:48-66
  std::vector<LqrKnot> knots;
  LqrKnot knb = generate_knot(nx, nu, nth, true);
  for (int i = 0; i < horz; i++) knots.push_back(knb);
  knots.push_back(generate_knot(nx, 0, nth, false));
  LqrProblem p(knots, nx);
  p.G0.setIdentity();
  for (int i = 0; i < nx; i++) p.g0[i] = -x0[i];
  return p;
}
static double maxabs(const std::array<double, 3> &e) { return std::max(e[0], std::max(e[1], e[2])); }
static double infdiff(const VectorOfVectors &a, const VectorOfVectors &b) {
  double m = 0; for (size_t i = 0; i < a.size(); i++) for (size_t j = 0; j < a[i].size(); j++) m = std::max(m, std::fabs(a[i][j] - b[i][j])); return m;
}

int main() {
  { // random_long_problem (riccati.cpp:113-135): nx=36, nu=12, N=100, mu = 1e-14, KKT <= 1e-9
    Vec x0(36); for (auto &x : x0) x = nrm(rng);
    LqrProblem prob = generate_problem(x0, 100, 36, 12);
    ProximalRiccatiSolver solver(prob);
    CHECK(solver.backward(1e-14, 1e-14), "backward");
    auto [xs, us, vs, lbdas] = lqrInitializeSolution(prob);
    CHECK(us.size() == 100 && xs.size() == 101, "solution sizes");
    CHECK(solver.forward(xs, us, vs, lbdas), "forward");
    const auto err = lqrComputeKktError(prob, xs, us, vs, lbdas, 0.0, 0.0);
    printf("random_long_problem KKT dyn %.2e cstr %.2e dual %.2e\n", err[0], err[1], err[2]);
    CHECK(maxabs(err) <= 1e-9, "KKT %.3e", maxabs(err));
    // feedback consistency: us[0] = kff + K xs[0] (riccati-kernel.hxx:317-322)
    const Vec ff = solver.getFeedforward(0), fb = solver.getFeedback(0);
    double worst = 0;
    for (int i = 0; i < 12; i++) { double u = ff[i]; for (int j = 0; j < 36; j++) u += fb[(size_t)i * 36 + j] * xs[0][j]; worst = std::max(worst, std::fabs(u - us[0][i])); }
    CHECK(worst <= 1e-9, "us[0] != kff + K x0: %.3e", worst);
  }
  { // one_knot_prob (riccati.cpp:93-111): horizon 0
    Vec x0(2, 1.0);
    LqrProblem prob = generate_problem(x0, 0, 2, 2);
    ProximalRiccatiSolver solver(prob);
    solver.backward(1e-13, 1e-13);
    auto [xs, us, vs, lbdas] = lqrInitializeSolution(prob);
    solver.forward(xs, us, vs, lbdas);
    const auto err = lqrComputeKktError(prob, xs, us, vs, lbdas, 0.0, 0.0);
    CHECK(maxabs(err) <= 1e-10, "one knot KKT %.3e", maxabs(err));
  }
  { // parametric (riccati.cpp:160-205): nx=3, nu=2, N=100, nth=1, mu=1e-12, KKT <= 1e-10
    Vec x0(3, 1.0);
    LqrProblem prob = generate_problem(x0, 100, 3, 2, 1);
    ProximalRiccatiSolver solver(prob);
    solver.backward(1e-12, 1e-12);
    auto [xs, us, vs, lbdas] = lqrInitializeSolution(prob);
    Vec theta{nrm(rng)};
    solver.forward(xs, us, vs, lbdas, theta);
    const auto err = lqrComputeKktError(prob, xs, us, vs, lbdas, 1e-12, 1e-12, theta);
    printf("parametric KKT %.2e; thGrad %.3e thHess %.3e\n", maxabs(err), solver.thGrad()[0], solver.thHess()[0]);
    CHECK(maxabs(err) <= 1e-10, "parametric KKT %.3e", maxabs(err));
    CHECK(std::isfinite(solver.thGrad()[0]) && std::isfinite(solver.thHess()[0]), "thGrad/thHess NaN");
  }
  { // parallel_solver_class (parallel.cpp:199-271): nx=32, nu=12, N=50, 6 legs, mu=1e-9
    Vec x0(32); for (auto &x : x0) x = nrm(rng);
    LqrProblem prob = generate_problem(x0, 50, 32, 12);
    ProximalRiccatiSolver ser(prob);
    ser.backward(1e-9, 1e-9);
    auto [xs_ref, us_ref, vs_ref, lb_ref] = lqrInitializeSolution(prob);
    ser.forward(xs_ref, us_ref, vs_ref, lb_ref);
    CHECK(maxabs(lqrComputeKktError(prob, xs_ref, us_ref, vs_ref, lb_ref, 1e-9, 1e-9)) <= 1e-10, "serial KKT");
    ParallelRiccatiSolver par(prob, 6);
    par.backward(1e-9, 1e-9);
    auto [xs, us, vs, lbdas] = lqrInitializeSolution(prob);
    par.forward(xs, us, vs, lbdas);
    const auto err = lqrComputeKktError(prob, xs, us, vs, lbdas, 1e-9, 1e-9);
    printf("parallel KKT %.2e, |xs - xs_ref| %.2e, |lbd - lbd_ref| %.2e\n", maxabs(err), infdiff(xs, xs_ref), infdiff(lbdas, lb_ref));
    CHECK(maxabs(err) <= 1e-10, "parallel KKT %.3e", maxabs(err));
    CHECK(infdiff(xs, xs_ref) <= 1e-10 && infdiff(lbdas, lb_ref) <= 1e-10, "serial != parallel");
    bool threw = false;
    try { ParallelRiccatiSolver bad(prob, 1); } catch (const std::runtime_error &) { threw = true; }
    CHECK(threw, "num_threads < 2 must throw (parallel-solver.hxx:25-29)");
  }
  { // random_long_problem, dense solver (riccati.cpp:143-157): same API, KKT <= 1e-9, same solution as the proximal one
    Vec x0(36); for (auto &x : x0) x = nrm(rng);
    LqrProblem prob = generate_problem(x0, 100, 36, 12);
    ProximalRiccatiSolver prox(prob);
    RiccatiSolverDense dense(prob);
    prox.backward(1e-14, 1e-14); dense.backward(1e-14, 1e-14);
    auto [xs, us, vs, lbdas] = lqrInitializeSolution(prob);
    auto [xd, ud, vd, ld] = lqrInitializeSolution(prob);
    prox.forward(xs, us, vs, lbdas); dense.forward(xd, ud, vd, ld);
    CHECK(maxabs(lqrComputeKktError(prob, xd, ud, vd, ld, 0.0, 0.0)) <= 1e-9, "dense KKT");
    CHECK(infdiff(xs, xd) <= 1e-10 && infdiff(us, ud) <= 1e-10, "dense != proximal");
    // the two are now DIFFERENT kernels (proximal sweep vs one Bunch-Kaufman of the stage KKT matrix per knot)
    double dmax = 0.0;
    for (int i = 0; i <= 100; i++) {
      const Vec a = prox.getFeedback((size_t)i), b = dense.getFeedback((size_t)i);
      for (size_t e = 0; e < a.size(); e++) dmax = std::fmax(dmax, std::fabs(a[e] - b[e]) / std::fmax(1.0, std::fabs(a[e])));
    }
    CHECK(dmax > 0.0 && dmax <= 1e-8, "dense and proximal gains: %g (must agree, yet come from different kernels)", dmax);
  }
  printf(fails ? "%d FAILURES\n" : "all C++ mirror tests passed\n", fails);
  return fails ? 1 : 0;
}
