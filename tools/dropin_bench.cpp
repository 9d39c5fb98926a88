// dropin_bench.cpp — end-to-end latency of the path the aligator ADAPTER takes:
//   backward(mudyn, mueq)   = garb200_stage_knot x (T+1) + stage_initial + commit (H2D) + kernels + sync
//   forward(xs, us, vs, lbdas) = kernel + D2H + scatter into std::vector<Vec>
//   updateGains()           = collapseFeedback + getFeedforward(i) / getFeedback(i) for i = 0 .. N
// i.e. SolverProxDDPTpl::innerLoop + updateGains (solver-proxddp.hxx:661-664, 316-334), through the C++
// mirror include/garb200.hpp (same call sequence as the adapter in include/garb200_aligator.hpp).
// Wall clock (std::chrono), host buffers, everything inside.  Prints one JSON object.
//   usage: dropin_bench [nx nu T reps]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include "garb200.hpp"

using namespace garb200;
using clk = std::chrono::steady_clock;
static std::mt19937 rng{42};
static std::normal_distribution<double> nrm;

static Mat wishart(int n, int p) {
  Mat W(n, p), out(n, n);
  for (auto &x : W.v) x = nrm(rng);
  for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) { double s = 0; for (int k = 0; k < p; k++) s += W(i, k) * W(j, k); out(i, j) = s; }
  return out;
}
static LqrKnot generate_knot(int nx, int nu, bool singular) {      // tests/gar/test_util.cpp:13-46
  LqrKnot k(nx, nu, 0, nx, 0);
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
  return k;
}

static double ms(clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); }

int main(int argc, char **argv) {
  const int nx = argc > 1 ? atoi(argv[1]) : 36, nu = argc > 2 ? atoi(argv[2]) : 12, T = argc > 3 ? atoi(argv[3]) : 100;
  const int reps = argc > 4 ? atoi(argv[4]) : 50;
  const double mu = 1e-11;                                           // bench/gar-riccati.cpp:26
  std::vector<LqrKnot> knots;
  for (int t = 0; t < T; t++) knots.push_back(generate_knot(nx, nu, true));   // distinct knots
  knots.push_back(generate_knot(nx, 0, false));
  LqrProblem prob(knots, nx);
  for (int i = 0; i < nx; i++) { prob.G0(i, i) = 1.0; prob.g0[i] = -nrm(rng); }
  ProximalRiccatiSolver solver(prob);
  auto [xs, us, vs, lbdas] = lqrInitializeSolution(prob);
  double tb = 0, tf = 0, tg = 0, sink = 0;
  for (int r = -3; r < reps; r++) {                                  // 3 warm-up rounds
    const auto t0 = clk::now();
    solver.backward(mu, mu);
    const auto t1 = clk::now();
    solver.forward(xs, us, vs, lbdas);
    const auto t2 = clk::now();
    solver.collapseFeedback();
    for (int i = 0; i <= T; i++) {                                   // updateGains (solver-proxddp.hxx:316-334)
      const Vec ff = solver.getFeedforward((size_t)i), fb = solver.getFeedback((size_t)i);
      sink += ff[0] + fb[0];
    }
    const auto t3 = clk::now();
    if (r >= 0) { tb += ms(t0, t1); tf += ms(t1, t2); tg += ms(t2, t3); }
  }
  const auto e = solver.kktError(mu, mu);
  printf("{\"workload\": \"1 x (nx=%d, nu=%d, T=%d) through the C++ mirror (adapter call sequence)\", \"reps\": %d, "
         "\"backward_ms\": %.4f, \"forward_ms\": %.4f, \"update_gains_ms\": %.4f, \"total_ms\": %.4f, "
         "\"kkt\": [%.3e, %.3e, %.3e], \"sink\": %.3e}\n",
         nx, nu, T, reps, tb / reps, tf / reps, tg / reps, (tb + tf + tg) / reps, e[0], e[1], e[2], sink);
  return 0;
}
