// garb200_aligator.hpp — the aligator-side adapter: a gar::RiccatiSolverBase<double> backed by the B200
// C ABI (include/garb200.h).  Drop this header into an aligator build (with -I <this repo>/include and
// -lgarb200) and construct `aligator::gar::B200RiccatiSolver` wherever ProximalRiccatiSolver /
// ParallelRiccatiSolver is constructed (solver-proxddp.hxx:163-187); nothing else changes for the caller.
//
// Interface implemented: /root/reference/include/aligator/gar/riccati-base.hpp:12-34
//   backward(mudyn, mueq)          :18      re-reads problem_->stages like proximal-riccati.hxx:37
//   forward(xs, us, vs, lbdas, th) :20-23   const, may be called repeatedly with different theta
//   cycleAppend(knot)              :25
//   collapseFeedback()             :29      (parallel-in-time mode only, parallel-solver.hpp:55-66)
//   getFeedforward(i) / getFeedback(i) :30-31  mutable Ref into solver-owned host mirrors, valid until the
//                                              next backward / cycleAppend; the first read after a
//                                              factorization mirrors ALL gains in one bulk D2H
//
// Only `.data()`, `.size()`, `.rows()`, `.cols()`, `.resize()` of the Eigen types are used, so that the
// header also compiles against the 40-line stub of riccati-base.hpp under tests/cpp/stub (CI in a container
// without Eigen: tests/test_cpp_mirror.py::test_adapter_header_compiles_against_stub).
#pragma once

#if !defined(__has_include)
#error "garb200_aligator.hpp needs __has_include (C++17)"
#elif !__has_include(<aligator/gar/riccati-base.hpp>)
#error "garb200_aligator.hpp: <aligator/gar/riccati-base.hpp> not found: add aligator's include directory (this header is the aligator-side adapter; the Eigen-free mirror is garb200.hpp)"
#endif

#include <aligator/gar/riccati-base.hpp>
#include <aligator/gar/lqr-problem.hpp>

#include <algorithm>
#include <cstring>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "garb200.h"

namespace aligator {
namespace gar {

class B200RiccatiSolver : public RiccatiSolverBase<double> {
public:
  using Base = RiccatiSolverBase<double>;
  using Scalar = double;
  using Knot = LqrKnotTpl<double>;
  using Problem = LqrProblemTpl<double>;
  using VectorXs = typename Base::VectorXs;
  using VectorRef = typename Base::VectorRef;
  using ConstVectorRef = typename Base::ConstVectorRef;
  using RowMatrixXs = typename Base::RowMatrixXs;
  using RowMatrixRef = typename Base::RowMatrixRef;

  /// num_legs = 0: ProximalRiccatiSolver semantics; >= 2: ParallelRiccatiSolver with that many legs
  /// (the problem is NOT mutated: the coupling blocks are read as views, parallel-solver.hpp:43-51).
  explicit B200RiccatiSolver(const Problem &problem, unsigned num_legs = 0, int device = 0)
      : problem_(&problem), legs_((int)num_legs) {
    if (num_legs == 1)   // ALIGATOR_RUNTIME_ERROR, parallel-solver.hxx:25-29
      throw std::runtime_error("(B200RiccatiSolver) num_legs should be 0 (serial) or >= 2");
    const int T = problem.horizon();
    if (T < 0) throw std::runtime_error("(B200RiccatiSolver) empty problem");
    nx_ = (int)problem.stages[0].nx;
    nu_t_.resize((size_t)T + 1);
    nc_t_.resize((size_t)T + 1);
    garb200_config cfg{};
    cfg.nx = nx_;
    cfg.nth = legs_ >= 2 ? 0 : (int)problem.ntheta();
    for (int t = 0; t <= T; t++) {
      const Knot &k = problem.stages[(size_t)t];
      if ((int)k.nx != nx_ || (int)k.nx2 != nx_) throw std::domain_error("(B200RiccatiSolver) nx2 == nx at every knot is required");
      nu_t_[(size_t)t] = (int)k.nu;
      nc_t_[(size_t)t] = (int)k.nc;
      cfg.nu = std::max(cfg.nu, (int)k.nu);
      cfg.nc = std::max(cfg.nc, (int)k.nc);
    }
    nu_ = cfg.nu; nc_ = cfg.nc; nth_ = cfg.nth; c0_ = (int)problem.nc0();
    cfg.horizon = T; cfg.nc0 = c0_; cfg.batch = 1; cfg.num_legs = legs_; cfg.device = device;
    cfg.flags = 0;
    cfg.nu_t = nu_t_.data(); cfg.nc_t = nc_t_.data();
    if (garb200_create(&cfg, &h_) != GARB200_OK || !h_)
      throw std::runtime_error(std::string("(B200RiccatiSolver) garb200_create: ") + garb200_last_error());
    const int nthg = legs_ >= 2 ? nx_ : nth_;      // fth width of the gains records
    ff_.resize((size_t)T + 1); fb_.resize((size_t)T + 1); fth_.resize((size_t)T + 1);
    for (int t = 0; t <= T; t++) {
      const int nr = nu_t_[(size_t)t] + nc_t_[(size_t)t] + 2 * nx_;
      ff_[(size_t)t].resize(nr);
      fb_[(size_t)t].resize(nr, nx_);
      fth_[(size_t)t].resize(nr, nthg);
    }
    X_.resize((size_t)(T + 1) * nx_);
    U_.resize((size_t)(T + 1) * std::max(nu_, 1));
    V_.resize((size_t)(T + 1) * std::max(nc_, 1));
    L_.resize((size_t)c0_ + (size_t)T * nx_);
  }
  ~B200RiccatiSolver() override { garb200_destroy(h_); }
  B200RiccatiSolver(const B200RiccatiSolver &) = delete;
  B200RiccatiSolver &operator=(const B200RiccatiSolver &) = delete;

  bool backward(const Scalar mudyn, const Scalar mueq) override {
    // the reference re-reads problem_->stages on every backward (proximal-riccati.hxx:37):
    // ManagedMatrix::data() pointers go straight into the view (managed-matrix.hpp:279-282)
    const Problem &p = *problem_;
    const int T = p.horizon();
    for (int t = 0; t <= T; t++) {
      const garb200_knot_view v = view_of(p.stages[(size_t)t]);
      if (garb200_stage_knot(h_, 0, t, &v) != GARB200_OK) return false;
    }
    if (garb200_stage_initial(h_, 0, p.G0.data(), p.g0.data()) != GARB200_OK) return false;
    if (garb200_commit(h_) != GARB200_OK) return false;
    gains_fresh_ = false;
    if (garb200_backward(h_, mudyn, mueq) != GARB200_OK) return false;
    return garb200_synchronize(h_) == GARB200_OK;      // GARB200_ERR_NUMERIC -> false
  }

  bool forward(std::vector<VectorXs> &xs, std::vector<VectorXs> &us, std::vector<VectorXs> &vs,
               std::vector<VectorXs> &lbdas,
               const std::optional<ConstVectorRef> &theta_ = std::nullopt) const override {
    const Problem &p = *problem_;
    const int T = p.horizon();
    const double *th = (theta_ && legs_ < 2 && nth_ > 0) ? theta_->data() : nullptr;   // parallel-solver.hxx:184
    const int rc = garb200_forward(h_, th, X_.data(), U_.data(), V_.data(), L_.data());
    if (rc != GARB200_OK && rc != GARB200_ERR_NUMERIC) return false;
    // packed trajectories (garb200.h header comment) -> the caller's pre-sized vectors (utils.hpp:139-168)
    for (int t = 0; t <= T; t++) {
      const size_t st = (size_t)t;
      const int m = nu_t_[st], c = nc_t_[st];
      std::memcpy(xs[st].data(), X_.data() + st * nx_, sizeof(double) * (size_t)nx_);
      if (st < us.size() && m > 0) std::memcpy(us[st].data(), U_.data() + st * nu_, sizeof(double) * (size_t)m);
      if (c > 0) std::memcpy(vs[st].data(), V_.data() + st * nc_, sizeof(double) * (size_t)c);
      if (t == 0) { if (c0_ > 0) std::memcpy(lbdas[0].data(), L_.data(), sizeof(double) * (size_t)c0_); }
      else std::memcpy(lbdas[st].data(), L_.data() + c0_ + (st - 1) * nx_, sizeof(double) * (size_t)nx_);
    }
    return rc == GARB200_OK;
  }

  void cycleAppend(const Knot &knot) override {
    const garb200_knot_view v = view_of(knot);
    if (garb200_cycle_append(h_, &v, /*shared=*/1) != GARB200_OK)
      throw std::runtime_error(std::string("(B200RiccatiSolver) cycleAppend: ") + garb200_last_error());
    gains_fresh_ = false;
  }

  void collapseFeedback() override {
    garb200_collapse_feedback(h_);
    gains_fresh_ = false;
  }

  VectorRef getFeedforward(size_t i) override { refresh(); return ff_[i]; }
  RowMatrixRef getFeedback(size_t i) override { refresh(); return fb_[i]; }
  /// datas[i].fth (riccati-kernel.hpp:78)
  RowMatrixRef getParamFeedback(size_t i) { refresh(); return fth_[i]; }

  /// lqrComputeKktError (utils.hxx:94-192) on the trajectory of the last forward(): {dyn, cstr, dual}
  bool kktError(Scalar mudyn, Scalar mueq, double out[3], const double *theta = nullptr) {
    return garb200_kkt_error(h_, mudyn, mueq, theta, out) == GARB200_OK;
  }
  garb200_solver *handle() { return h_; }

private:
  static garb200_knot_view view_of(const Knot &k) {
    garb200_knot_view v;
    v.nu = (int)k.nu; v.nc = (int)k.nc;
    v.Q = k.Q.data(); v.S = k.S.data(); v.R = k.R.data(); v.q = k.q.data(); v.r = k.r.data();
    v.A = k.A.data(); v.B = k.B.data(); v.E = k.E.data(); v.f = k.f.data();
    v.C = k.C.data(); v.D = k.D.data(); v.d = k.d.data();
    v.Gth = k.Gth.data(); v.Gx = k.Gx.data(); v.Gu = k.Gu.data(); v.gamma = k.gamma.data();
    return v;
  }
  /// updateGains() reads every knot right after backward (solver-proxddp.hxx:316-334): mirror them all at
  /// the first read (the C ABI does ONE bulk D2H and serves the rest from pinned memory); row-major fb on
  /// the device == Eigen RowMajor, so each mirror is a plain copy.
  void refresh() {
    if (gains_fresh_) return;
    const int T = problem_->horizon();
    for (int t = 0; t <= T; t++) {
      const size_t st = (size_t)t;
      garb200_get_gains(h_, 0, t, ff_[st].data(), fb_[st].data(), fth_[st].size() ? fth_[st].data() : nullptr);
    }
    gains_fresh_ = true;
  }

  const Problem *problem_;
  garb200_solver *h_ = nullptr;
  int legs_ = 0, nx_ = 0, nu_ = 0, nc_ = 0, nth_ = 0, c0_ = 0;
  std::vector<int> nu_t_, nc_t_;
  std::vector<VectorXs> ff_;
  std::vector<RowMatrixXs> fb_, fth_;
  mutable std::vector<double> X_, U_, V_, L_;
  bool gains_fresh_ = false;
};

} // namespace gar
} // namespace aligator
