// TEST STUB of <aligator/gar/lqr-problem.hpp>: the data members and accessors of LqrKnotTpl / LqrProblemTpl
// that the adapter reads (lqr-problem.hpp:32-56, 198-264 of the reference), std::vector-backed.
#pragma once
#include "riccati-base.hpp"

namespace aligator {
namespace gar {
using uint = unsigned int;

template <typename Scalar> struct LqrKnotTpl {
  uint nx = 0, nu = 0, nc = 0, nx2 = 0, nth = 0;
  StubMatrix<false> Q, S, R, A, B, E, C, D, Gth, Gx, Gu, Gv;
  StubVector q, r, f, d, gamma;
  LqrKnotTpl() = default;
  LqrKnotTpl(uint nx_, uint nu_, uint nc_, uint nth_ = 0)
      : nx(nx_), nu(nu_), nc(nc_), nx2(nx_), nth(nth_), Q(nx_, nx_), S(nx_, nu_), R(nu_, nu_), A(nx_, nx_), B(nx_, nu_),
        E(nx_, nx_), C(nc_, nx_), D(nc_, nu_), Gth(nth_, nth_), Gx(nx_, nth_), Gu(nu_, nth_), Gv(nc_, nth_), q(nx_), r(nu_),
        f(nx_), d(nc_), gamma(nth_) {}
};

template <typename Scalar> struct LqrProblemTpl {
  StubMatrix<false> G0;
  StubVector g0;
  std::vector<LqrKnotTpl<Scalar>> stages;
  LqrProblemTpl(const std::vector<LqrKnotTpl<Scalar>> &knots, long nc0)
      : G0(nc0, knots.empty() ? 0 : (long)knots[0].nx), g0(nc0), stages(knots) {}
  int horizon() const noexcept { return (int)stages.size() - 1; }
  uint nc0() const noexcept { return (uint)g0.rows(); }
  uint ntheta() const { return stages[0].nth; }
};

} // namespace gar
} // namespace aligator
