// garb200.hpp — header-only, Eigen-free C++17 mirror of the reference's `gar` interface over
// the C ABI (garb200.h).  Same names, argument meaning and error behaviour as
//   LqrKnotTpl / LqrProblemTpl        /root/reference/include/aligator/gar/lqr-problem.hpp:32-289
//   RiccatiSolverBase                 gar/riccati-base.hpp:12-34
//   ProximalRiccatiSolver             gar/proximal-riccati.hpp:12-45
//   ParallelRiccatiSolver             gar/parallel-solver.hpp:24-108
//   lqrInitializeSolution             gar/utils.hpp:139-168
//   lqrComputeKktError                gar/utils.hxx:94-192
// Matrices are plain column-major std::vector<double> (what `ManagedMatrix::data()` exposes,
// memory/managed-matrix.hpp:279-282).  Nothing here computes on the CPU.
#pragma once
#include <algorithm>
#include <array>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>
#include "garb200.h"

namespace garb200 {

using Vec = std::vector<double>;

/// Column-major dense matrix (rows x cols); `data()` is what the C ABI consumes.
struct Mat {
  int rows = 0, cols = 0;
  Vec v;
  Mat() = default;
  Mat(int r, int c) : rows(r), cols(c), v((size_t)r * c, 0.0) {}
  double &operator()(int i, int j) { return v[(size_t)i + (size_t)j * rows]; }
  double operator()(int i, int j) const { return v[(size_t)i + (size_t)j * rows]; }
  const double *data() const { return v.empty() ? nullptr : v.data(); }
  double *data() { return v.empty() ? nullptr : v.data(); }
  void setZero() { std::fill(v.begin(), v.end(), 0.0); }
  void setIdentity(double s = 1.0) { setZero(); for (int i = 0; i < std::min(rows, cols); i++) (*this)(i, i) = s; }
};

/// lqr-problem.hpp:32-196 — cost 1/2 [x;u]^T [Q S; S^T R] [x;u] + q^T x + r^T u,
/// dynamics E x' + A x + B u + f = 0, constraint C x + D u + d = 0, parameter block G*.
struct LqrKnot {
  int nx = 0, nu = 0, nc = 0, nx2 = 0, nth = 0;
  Mat Q, S, R, A, B, E, C, D, Gth, Gx, Gu, Gv;
  Vec q, r, f, d, gamma;
  LqrKnot() = default;
  LqrKnot(int nx_, int nu_, int nc_, int nx2_ = -1, int nth_ = 0)
      : nx(nx_), nu(nu_), nc(nc_), nx2(nx2_ < 0 ? nx_ : nx2_), nth(nth_), Q(nx_, nx_), S(nx_, nu_),
        R(nu_, nu_), A(nx2, nx_), B(nx2, nu_), E(nx2, nx2), C(nc_, nx_), D(nc_, nu_), q(nx_, 0.0),
        r(nu_, 0.0), f(nx2, 0.0), d(nc_, 0.0) {
    if (nx2 != nx) throw std::domain_error("garb200: nx2 == nx required");
    addParameterization(nth_);
  }
  /// lqr-problem.hxx:215-223
  void addParameterization(int nth_) {
    nth = nth_;
    Gth = Mat(nth, nth); Gx = Mat(nx, nth); Gu = Mat(nu, nth); Gv = Mat(nc, nth);
    gamma.assign(nth, 0.0);
  }
  garb200_knot_view view() const {
    garb200_knot_view k{};
    k.nu = nu; k.nc = nc;
    k.Q = Q.data(); k.S = S.data(); k.R = R.data(); k.q = q.data(); k.r = r.data();
    k.A = A.data(); k.B = B.data(); k.E = E.data(); k.f = f.data();
    k.C = C.data(); k.D = D.data(); k.d = d.data();
    k.Gth = Gth.data(); k.Gx = Gx.data(); k.Gu = Gu.data(); k.gamma = gamma.data();
    return k;
  }
};

/// lqr-problem.hpp:198-289
struct LqrProblem {
  std::vector<LqrKnot> stages;
  Mat G0;
  Vec g0;
  LqrProblem() = default;
  LqrProblem(const std::vector<LqrKnot> &knots, int nc0)
      : stages(knots), G0(nc0, knots.empty() ? 0 : knots[0].nx), g0(nc0, 0.0) {}
  int horizon() const { return (int)stages.size() - 1; }
  int nc0() const { return (int)g0.size(); }
  int ntheta() const { return stages.empty() ? 0 : stages[0].nth; }
  bool isInitialized() const { return !stages.empty(); }
  bool isParameterized() const { return isInitialized() && ntheta() > 0; }
  void addParameterization(int nth) { for (auto &k : stages) k.addParameterization(nth); }
};

using VectorOfVectors = std::vector<Vec>;

/// utils.hpp:139-168: zeroed xs, us, vs, lbdas of the right sizes (us drops the terminal
/// entry when the terminal knot has nu == 0).
inline std::array<VectorOfVectors, 4> lqrInitializeSolution(const LqrProblem &p) {
  const int N = p.horizon();
  VectorOfVectors xs(N + 1), us(N + 1), vs(N + 1), lbdas(N + 1);
  lbdas[0].assign(p.nc0(), 0.0);
  for (int i = 0; i <= N; i++) {
    const LqrKnot &k = p.stages[i];
    xs[i].assign(k.nx, 0.0); us[i].assign(k.nu, 0.0); vs[i].assign(k.nc, 0.0);
    if (i < N) lbdas[i + 1].assign(k.nx2, 0.0);
  }
  if (p.stages.back().nu == 0) us.pop_back();
  return {xs, us, vs, lbdas};
}

/// riccati-base.hpp:12-34 — the drop-in boundary.
class RiccatiSolverBase {
public:
  virtual bool backward(double mudyn, double mueq) = 0;
  virtual bool forward(VectorOfVectors &xs, VectorOfVectors &us, VectorOfVectors &vs,
                       VectorOfVectors &lbdas, const std::optional<Vec> &theta = std::nullopt) const = 0;
  virtual void cycleAppend(const LqrKnot &knot) = 0;
  virtual void collapseFeedback() {}
  virtual Vec getFeedforward(size_t i) = 0;
  /// row-major (nu + nc + 2 nx) x nx, like RowMatrixRef
  virtual Vec getFeedback(size_t i) = 0;
  virtual ~RiccatiSolverBase() = default;
};

namespace detail {
inline void check(int rc, const char *what) {
  if (rc != GARB200_OK && rc != GARB200_ERR_NUMERIC)
    throw std::runtime_error(std::string(what) + ": " + garb200_last_error());
}

/// Shared implementation: one problem (batch = 1) behind a garb200_solver handle.
class GpuSolver : public RiccatiSolverBase {
public:
  GpuSolver(const LqrProblem &problem, int num_legs, int device, int extra_flags = 0) : problem_(&problem) {
    const int T = problem.horizon();
    if (T < 0) throw std::runtime_error("garb200: empty problem");
    nx_ = problem.stages[0].nx;
    nu_t_.resize(T + 1); nc_t_.resize(T + 1);
    garb200_config cfg{};
    cfg.nx = nx_; cfg.nth = num_legs >= 2 ? 0 : problem.ntheta();
    for (int t = 0; t <= T; t++) {
      const LqrKnot &k = problem.stages[t];
      if (k.nx != nx_) throw std::domain_error("garb200: the same nx at every knot is required");
      nu_t_[t] = k.nu; nc_t_[t] = k.nc;
      cfg.nu = std::max(cfg.nu, k.nu); cfg.nc = std::max(cfg.nc, k.nc);
    }
    nu_ = cfg.nu; nc_ = cfg.nc; nth_ = cfg.nth; legs_ = num_legs;
    cfg.horizon = T; cfg.nc0 = problem.nc0(); cfg.batch = 1; cfg.num_legs = num_legs;
    cfg.device = device;
    // datas[t].vm of every knot belongs to the serial solver's public state; the parallel-in-time solver
    // keeps the leg heads only (storing all of them would take it off the specialised leg sweep)
    cfg.flags = (num_legs >= 2 ? 0 : GARB200_STORE_ALL_VALUES) | extra_flags;
    cfg.nu_t = nu_t_.data(); cfg.nc_t = nc_t_.data();
    check(garb200_create(&cfg, &h_), "garb200_create");
    if (!h_) throw std::runtime_error(std::string("garb200_create: ") + garb200_last_error());
    X_.resize((size_t)(T + 1) * nx_); U_.resize((size_t)(T + 1) * std::max(nu_, 1));
    V_.resize((size_t)(T + 1) * std::max(nc_, 1)); L_.resize((size_t)problem.nc0() + (size_t)T * nx_);
  }
  ~GpuSolver() override { garb200_destroy(h_); }
  GpuSolver(const GpuSolver &) = delete;
  GpuSolver &operator=(const GpuSolver &) = delete;

  /// The reference re-reads problem_->stages on every backward() (proximal-riccati.hxx:37).
  bool backward(double mudyn, double mueq) override {
    const LqrProblem &p = *problem_;
    for (int t = 0; t <= p.horizon(); t++) {
      const garb200_knot_view v = p.stages[t].view();
      check(garb200_stage_knot(h_, 0, t, &v), "garb200_stage_knot");
    }
    check(garb200_stage_initial(h_, 0, p.G0.data(), p.g0.data()), "garb200_stage_initial");
    check(garb200_commit(h_), "garb200_commit");
    check(garb200_backward(h_, mudyn, mueq), "garb200_backward");
    const int rc = garb200_synchronize(h_);
    check(rc, "garb200_synchronize");
    return rc == GARB200_OK;
  }

  bool forward(VectorOfVectors &xs, VectorOfVectors &us, VectorOfVectors &vs, VectorOfVectors &lbdas,
               const std::optional<Vec> &theta = std::nullopt) const override {
    const LqrProblem &p = *problem_;
    const int T = p.horizon(), c0 = p.nc0();
    const double *th = (theta && legs_ < 2 && nth_ > 0) ? theta->data() : nullptr;   // parallel-solver.hxx:184
    const int rc = garb200_forward(h_, th, X_.data(), U_.data(), V_.data(), L_.data());
    check(rc, "garb200_forward");
    for (int t = 0; t <= T; t++) {
      const LqrKnot &k = p.stages[t];
      xs[t].assign(X_.begin() + (size_t)t * nx_, X_.begin() + (size_t)(t + 1) * nx_);
      if (t < (int)us.size() && k.nu > 0) us[t].assign(U_.begin() + (size_t)t * nu_, U_.begin() + (size_t)t * nu_ + k.nu);
      vs[t].assign(V_.begin() + (size_t)t * nc_, V_.begin() + (size_t)t * nc_ + k.nc);
      if (t == 0) lbdas[0].assign(L_.begin(), L_.begin() + c0);
      else lbdas[t].assign(L_.begin() + c0 + (size_t)(t - 1) * nx_, L_.begin() + c0 + (size_t)t * nx_);
    }
    return rc == GARB200_OK;
  }

  void cycleAppend(const LqrKnot &knot) override {
    const garb200_knot_view v = knot.view();
    check(garb200_cycle_append(h_, &v, 1), "garb200_cycle_append");
  }
  Vec getFeedforward(size_t i) override {
    Vec ff(rows(i));
    check(garb200_get_gains(h_, 0, (int)i, ff.data(), nullptr, nullptr), "garb200_get_gains");
    return ff;
  }
  Vec getFeedback(size_t i) override {
    Vec fb((size_t)rows(i) * nx_);
    check(garb200_get_gains(h_, 0, (int)i, nullptr, fb.data(), nullptr), "garb200_get_gains");
    return fb;
  }
  /// lqrComputeKktError on the trajectory of the last forward(): {dyn, cstr, dual}
  std::array<double, 3> kktError(double mudyn, double mueq, const std::optional<Vec> &theta = std::nullopt) {
    std::array<double, 3> out{};
    check(garb200_kkt_error(h_, mudyn, mueq, theta ? theta->data() : nullptr, out.data()), "garb200_kkt_error");
    return out;
  }
  garb200_solver *handle() { return h_; }

protected:
  int rows(size_t i) const { return nu_t_[i] + nc_t_[i] + 2 * nx_; }
  const LqrProblem *problem_;
  garb200_solver *h_ = nullptr;
  int nx_ = 0, nu_ = 0, nc_ = 0, nth_ = 0, legs_ = 0;
  std::vector<int> nu_t_, nc_t_;
  mutable Vec X_, U_, V_, L_;
};
} // namespace detail

/// proximal-riccati.hpp:12-45
class ProximalRiccatiSolver : public detail::GpuSolver {
public:
  explicit ProximalRiccatiSolver(const LqrProblem &problem, int device = 0, int extra_flags = 0)
      : GpuSolver(problem, 0, device, extra_flags) {}
  struct kkt0_t { Vec ff, fth; };
  kkt0_t kkt0() {
    const int s0 = nx_ + problem_->nc0();
    kkt0_t k{Vec(s0), Vec((size_t)s0 * nth_)};
    detail::check(garb200_get_initial(h_, 0, k.ff.data(), k.fth.data(), nullptr, nullptr), "garb200_get_initial");
    return k;
  }
  Vec thGrad() { Vec g(nth_); detail::check(garb200_get_initial(h_, 0, nullptr, nullptr, g.data(), nullptr), "garb200_get_initial"); return g; }
  Vec thHess() { Vec H((size_t)nth_ * nth_); detail::check(garb200_get_initial(h_, 0, nullptr, nullptr, nullptr, H.data()), "garb200_get_initial"); return H; }
};

/// dense-riccati.hpp:28-58 — same constructor and virtuals, GPU-backed by csrc/dense.cu: one Bunch-Kaufman
/// factorization of the (nu + nc + 2 nx)^2 stage KKT matrix per knot (dense-kernel.hpp:56-212).  Reads Gv
/// (dense-kernel.hpp:76,134), which the proximal kernel ignores (riccati-kernel.hxx:128,272).
class RiccatiSolverDense : public ProximalRiccatiSolver {
public:
  explicit RiccatiSolverDense(const LqrProblem &problem, int device = 0)
      : ProximalRiccatiSolver(problem, device, GARB200_DENSE_KERNEL) {}
  bool backward(double mudyn, double mueq) override {
    const LqrProblem &p = *problem_;
    const int T = p.horizon();
    if (nc_ * nth_ > 0) {                        // Gv is not part of the packed knot record
      gv_.assign((size_t)(T + 1) * nc_ * nth_, 0.0);
      for (int t = 0; t <= T; t++) {
        const LqrKnot &k = p.stages[t];
        for (int j = 0; j < nth_; j++)
          for (int i = 0; i < k.nc; i++) gv_[(size_t)t * nc_ * nth_ + i + (size_t)j * nc_] = k.Gv(i, j);
      }
      detail::check(garb200_upload_gv(h_, gv_.data()), "garb200_upload_gv");
    }
    return ProximalRiccatiSolver::backward(mudyn, mueq);
  }

private:
  Vec gv_;
};

/// parallel-solver.hpp:24-108 — legs are CTAs instead of OpenMP threads.
class ParallelRiccatiSolver : public detail::GpuSolver {
public:
  ParallelRiccatiSolver(LqrProblem &problem, unsigned num_threads, int device = 0)
      : GpuSolver(check_threads(problem, num_threads), (int)num_threads, device), numThreads(num_threads) {}
  void collapseFeedback() override { detail::check(garb200_collapse_feedback(h_), "garb200_collapse_feedback"); }
  unsigned numThreads;

private:
  static const LqrProblem &check_threads(const LqrProblem &p, unsigned n) {
    if (n < 2)   // ALIGATOR_RUNTIME_ERROR, parallel-solver.hxx:25-29
      throw std::runtime_error("(ParallelRiccatiSolver) num_threads (=" + std::to_string(n) +
                               ") should be greater than or equal to 2.");
    return p;
  }
};

/// utils.hxx:94-192 as a free function, like the reference: scores ANY trajectory.
inline std::array<double, 3> lqrComputeKktError(const LqrProblem &problem, const VectorOfVectors &xs,
                                                const VectorOfVectors &us, const VectorOfVectors &vs,
                                                const VectorOfVectors &lbdas, double mudyn, double mueq,
                                                const std::optional<Vec> &theta = std::nullopt, int device = 0) {
  ProximalRiccatiSolver s(problem, device);
  const int T = problem.horizon(), nx = problem.stages[0].nx, c0 = problem.nc0();
  int nu = 0, nc = 0;
  for (auto &k : problem.stages) { nu = std::max(nu, k.nu); nc = std::max(nc, k.nc); }
  Vec X((size_t)(T + 1) * nx, 0.0), U((size_t)(T + 1) * std::max(nu, 1), 0.0), V((size_t)(T + 1) * std::max(nc, 1), 0.0),
      L((size_t)c0 + (size_t)T * nx, 0.0);
  for (int t = 0; t <= T; t++) {
    std::copy(xs[t].begin(), xs[t].end(), X.begin() + (size_t)t * nx);
    if (t < (int)us.size()) std::copy(us[t].begin(), us[t].end(), U.begin() + (size_t)t * nu);
    std::copy(vs[t].begin(), vs[t].end(), V.begin() + (size_t)t * nc);
    if (t == 0) std::copy(lbdas[0].begin(), lbdas[0].end(), L.begin());
    else std::copy(lbdas[t].begin(), lbdas[t].end(), L.begin() + c0 + (size_t)(t - 1) * nx);
  }
  garb200_solver *h = s.handle();
  for (int t = 0; t <= T; t++) { const garb200_knot_view v = problem.stages[t].view(); detail::check(garb200_stage_knot(h, 0, t, &v), "stage_knot"); }
  detail::check(garb200_stage_initial(h, 0, problem.G0.data(), problem.g0.data()), "stage_initial");
  detail::check(garb200_commit(h), "commit");
  detail::check(garb200_upload_trajectory(h, X.data(), U.data(), V.data(), L.data()), "upload_trajectory");
  std::array<double, 3> out{};
  detail::check(garb200_kkt_error(h, mudyn, mueq, theta ? theta->data() : nullptr, out.data()), "kkt_error");
  return out;
}

} // namespace garb200
