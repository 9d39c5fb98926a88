// TEST STUB of <aligator/gar/riccati-base.hpp> (+ the typedefs of aligator/math.hpp it pulls in) for a
// container without Eigen: std::vector-backed stand-ins with the handful of members the adapter
// include/garb200_aligator.hpp uses (.data(), .size(), .rows(), .cols(), .resize()).  It exists so that the
// adapter's syntax and its override signatures (riccati-base.hpp:18-31 of the reference) are checked by CI
// here; it is NOT aligator.
#pragma once
#include <cstddef>
#include <optional>
#include <vector>

namespace aligator {
namespace gar {

struct StubVector {
  std::vector<double> v;
  StubVector() = default;
  explicit StubVector(long n) : v((size_t)n, 0.0) {}
  double *data() { return v.data(); }
  const double *data() const { return v.data(); }
  long size() const { return (long)v.size(); }
  long rows() const { return (long)v.size(); }
  void resize(long n) { v.assign((size_t)n, 0.0); }
  double &operator[](long i) { return v[(size_t)i]; }
  double operator[](long i) const { return v[(size_t)i]; }
};
template <bool RowMajor> struct StubMatrix {
  std::vector<double> v;
  long r = 0, c = 0;
  StubMatrix() = default;
  StubMatrix(long rows, long cols) { resize(rows, cols); }
  double *data() { return v.data(); }
  const double *data() const { return v.data(); }
  long size() const { return r * c; }
  long rows() const { return r; }
  long cols() const { return c; }
  void resize(long rows, long cols) { r = rows; c = cols; v.assign((size_t)(rows * cols), 0.0); }
  double &operator()(long i, long j) { return v[(size_t)(RowMajor ? i * c + j : i + j * r)]; }
  double operator()(long i, long j) const { return v[(size_t)(RowMajor ? i * c + j : i + j * r)]; }
};
template <class M> struct StubRef {            // Eigen::Ref<M>: a mutable view into someone else's storage
  double *p; long n, r, c;
  StubRef(M &m) : p(m.data()), n(m.size()), r(m.rows()), c(m.size() ? m.size() / m.rows() : 0) {}
  double *data() const { return p; }
  long size() const { return n; }
  long rows() const { return r; }
  long cols() const { return c; }
};
struct StubConstVectorRef {                    // Eigen::Ref<const VectorXs>
  const double *p; long n;
  StubConstVectorRef(const StubVector &m) : p(m.data()), n(m.size()) {}
  const double *data() const { return p; }
  long size() const { return n; }
};

template <typename Scalar> struct LqrKnotTpl;
template <typename Scalar> struct LqrProblemTpl;

template <typename _Scalar> class RiccatiSolverBase {
public:
  using Scalar = _Scalar;
  using LqrKnot = LqrKnotTpl<Scalar>;
  using VectorXs = StubVector;
  using MatrixXs = StubMatrix<false>;
  using VectorRef = StubRef<StubVector>;
  using ConstVectorRef = StubConstVectorRef;
  using RowMatrixXs = StubMatrix<true>;
  using RowMatrixRef = StubRef<StubMatrix<true>>;

  virtual bool backward(const Scalar mudyn, const Scalar mueq) = 0;
  virtual bool forward(std::vector<VectorXs> &xs, std::vector<VectorXs> &us, std::vector<VectorXs> &vs,
                       std::vector<VectorXs> &lbdas,
                       const std::optional<ConstVectorRef> &theta_ = std::nullopt) const = 0;
  virtual void cycleAppend(const LqrKnot &knot) = 0;
  virtual void collapseFeedback() {}
  virtual VectorRef getFeedforward(size_t) = 0;
  virtual RowMatrixRef getFeedback(size_t) = 0;
  virtual ~RiccatiSolverBase() = default;
};

} // namespace gar
} // namespace aligator
