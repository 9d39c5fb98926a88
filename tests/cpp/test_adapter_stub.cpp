// include/garb200_aligator.hpp (the aligator-side adapter) compiled against the stub headers of
// tests/cpp/stub and driven through the RiccatiSolverBase interface: backward / forward / getFeedback /
// getFeedforward / cycleAppend / collapseFeedback.  Exit code 0 = all assertions hold (needs a CUDA device).
#include <cmath>
#include <cstdio>
#include <random>
#include "garb200_aligator.hpp"

using namespace aligator::gar;
using Knot = LqrKnotTpl<double>;
using Problem = LqrProblemTpl<double>;
using Vec = RiccatiSolverBase<double>::VectorXs;
static std::mt19937 rng{7};
static std::normal_distribution<double> nrm;
static int fails = 0;
#define CHECK(cond, ...) do { if (!(cond)) { fails++; printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } } while (0)

static Knot make_knot(unsigned nx, unsigned nu) {
  Knot k(nx, nu, 0);
  const unsigned d = nx + nu, p = d + 1;
  std::vector<double> W(d * p);
  for (auto &x : W) x = nrm(rng);
  auto w = [&](unsigned i, unsigned j) { double s = 0; for (unsigned l = 0; l < p; l++) s += W[i + l * d] * W[j + l * d]; return s; };
  for (unsigned i = 0; i < nx; i++) for (unsigned j = 0; j < nx; j++) k.Q(i, j) = w(i, j);
  for (unsigned i = 0; i < nx; i++) for (unsigned j = 0; j < nu; j++) k.S(i, j) = w(i, nx + j);
  for (unsigned i = 0; i < nu; i++) for (unsigned j = 0; j < nu; j++) k.R(i, j) = w(nx + i, nx + j);
  for (auto &x : k.q.v) x = nrm(rng);
  for (auto &x : k.r.v) x = nrm(rng);
  for (auto &x : k.A.v) x = 0.3 * nrm(rng);
  for (auto &x : k.B.v) x = nrm(rng);
  for (unsigned i = 0; i < nx; i++) k.E(i, i) = -1.0;
  for (auto &x : k.f.v) x = nrm(rng);
  return k;
}

static void run(unsigned nx, unsigned nu, int T, unsigned legs) {
  std::vector<Knot> knots;
  for (int t = 0; t < T; t++) knots.push_back(make_knot(nx, nu));
  knots.push_back(make_knot(nx, 0));
  Problem prob(knots, nx);
  for (unsigned i = 0; i < nx; i++) { prob.G0(i, i) = 1.0; prob.g0[i] = nrm(rng); }
  B200RiccatiSolver solver(prob, legs);
  RiccatiSolverBase<double> &base = solver;        // everything below goes through the reference interface
  std::vector<Vec> xs, us, vs, lbdas;
  for (int t = 0; t <= T; t++) { xs.emplace_back(nx); vs.emplace_back(0); if (t < T) us.emplace_back(nu); lbdas.emplace_back(nx); }
  const double mu = 1e-9;
  CHECK(base.backward(mu, mu), "backward");
  CHECK(base.forward(xs, us, vs, lbdas), "forward");
  double e[3];
  CHECK(solver.kktError(mu, mu, e) && std::fmax(e[0], std::fmax(e[1], e[2])) <= 1e-8, "kkt %g %g %g", e[0], e[1], e[2]);
  base.collapseFeedback();
  auto K = base.getFeedback(0);
  auto kff = base.getFeedforward(0);
  CHECK(K.rows() == (long)(nu + 2 * nx) && K.cols() == (long)nx && kff.size() == (long)(nu + 2 * nx), "gain shapes");
  if (legs < 2)                                    // u0 = kff + K x0 (riccati-kernel.hxx:317-322)
    for (unsigned i = 0; i < nu; i++) {
      double u = kff.data()[i];
      for (unsigned j = 0; j < nx; j++) u += K.data()[i * nx + j] * xs[0][j];
      CHECK(std::fabs(u - us[0][i]) <= 1e-10 * std::fmax(1.0, std::fabs(u)), "u0[%u] %g vs %g", i, u, us[0][i]);
    }
  // a mutable Ref into solver-owned memory, stable across calls until the next backward
  K.data()[0] = 123.0;
  CHECK(base.getFeedback(0).data()[0] == 123.0, "getFeedback must return a view of solver-owned memory");
  // the solver re-reads the problem on every backward (proximal-riccati.hxx:37)
  const double u00 = us[0][0];
  for (auto &x : prob.stages[1].q.v) x = 5.0;
  CHECK(base.backward(mu, mu) && base.forward(xs, us, vs, lbdas), "re-solve");
  CHECK(std::fabs(us[0][0] - u00) > 1e-9, "problem change not seen");
  CHECK(base.getFeedback(0).data()[0] != 123.0, "gains mirror must refresh after backward");
  if (legs < 2) {
    base.cycleAppend(prob.stages[0]);
    CHECK(base.backward(mu, mu), "backward after cycleAppend");
  }
}

int main() {
  run(4, 2, 6, 0);
  run(36, 12, 20, 0);
  run(12, 4, 30, 3);
  if (fails == 0) printf("all adapter (stub) tests passed\n");
  return fails == 0 ? 0 : 1;
}
