// oracle/gar_oracle.hpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// CPU restatement (plain C++17, no Eigen) of aligator's `gar` LQ sub-problem solver,
// used as the parity oracle for the CUDA path and as the timed CPU baseline ("port").
// Every function cites the reference file:line it follows (paths relative to
// /root/reference/include/aligator/gar unless noted).
//
// PARITY PINNING.  The reference holds NO golden vectors for this path and cannot be
// compiled here (needs Eigen/Boost, absent).  The oracle is pinned at the
// SOLVE level the way the reference pins itself: its own test assertions
// (tests/gar/riccati.cpp, tests/gar/parallel.cpp: KKT residual <= 1e-9 / 1e-10,
// serial == parallel to 1e-10) re-run on the oracle, plus agreement with an independent
// dense LAPACK solve of the `lqrDenseMatrix` KKT system.  At the level of individual
// Eigen routines parity is UNPINNED (rounding differs with Eigen's blocking).
#pragma once
#include "bunchkaufman.hpp"
#include "linalg.hpp"
#include <array>
#include <optional>

namespace gar_oracle {

/// lqr-problem.hpp:32-196 (LqrKnotTpl); sizes lqr-problem.hxx:38-54.
struct Knot {
  int nx = 0, nu = 0, nc = 0, nx2 = 0, nth = 0;
  Mat Q, S, R, q, r;
  Mat A, B, E, f;
  Mat C, D, d;
  Mat Gth, Gx, Gu, Gv, gamma;
  Knot() = default;
  Knot(int nx, int nu, int nc, int nth = 0);
  /// lqr-problem.hxx:215-223: re-allocate the parameter block with zeros.
  void addParameterization(int nth);
};

/// lqr-problem.hpp:198-289 (LqrProblemTpl).
struct Problem {
  Mat G0, g0;
  std::vector<Knot> stages;
  int horizon() const { return (int)stages.size() - 1; }
  int nc0() const { return g0.r; }
  int ntheta() const { return stages.empty() ? 0 : stages[0].nth; }
  void addParameterization(int nth) { for (auto &k : stages) k.addParameterization(nth); }
};

/// riccati-kernel.hpp:39-60 (StageFactor::value_t)
struct Value {
  Mat Pmat, pvec, Vxx, vx, Vxt, Vtt, vt;
  Value() = default;
  Value(int nx, int nth);
};

/// riccati-kernel.hpp:35-91 (StageFactor). ff/fb/fth hold block rows [K;Z;L;A]
/// (sizes [nu;nc;nx;nx]); they are kept column-major here and exported row-major.
struct StageFactor {
  int nx = 0, nu = 0, nc = 0, nth = 0;
  Mat ff, fb, fth;
  Value vm;
  // scratch kept for inspection (riccati-kernel.hpp:64-90)
  Mat Qhat, Rhat, Shat, qhat, rhat, AtV, BtV, Guhat, Einv, Ptilde;
  PartialPivLU Efact;
  LLT schurChol;
  BunchKaufman kktChol;
  StageFactor() = default;
  StageFactor(int nx, int nu, int nc, int nth);
};

/// riccati-kernel.hpp:101-110 (kkt0_t)
struct Kkt0 {
  Mat ff, fth; // rows [x0 (nx); lbd0 (nc0)]
  BunchKaufman chol;
};

/// riccati-kernel.hxx:90-150
void terminalSolve(const Knot &model, double mueq, StageFactor &d);
/// riccati-kernel.hxx:167-300
void stageKernelSolve(const Knot &model, StageFactor &d, Value &vn, double mudyn, double mueq);
/// riccati-kernel.hxx:64-88 on the span [beg, end)
bool backwardImpl(const std::vector<Knot> &stages, int beg, int end, double mudyn,
                  double mueq, std::vector<StageFactor> &datas);
/// riccati-kernel.hxx:303-361 on the span [beg, end); xs/us/vs/lbdas are full-horizon
bool forwardImpl(const std::vector<Knot> &stages, const std::vector<StageFactor> &datas,
                 int beg, int end, std::vector<Mat> &xs, std::vector<Mat> &us,
                 std::vector<Mat> &vs, std::vector<Mat> &lbdas, const Mat *theta);

/// proximal-riccati.hpp:12-45 / proximal-riccati.hxx
struct ProximalRiccatiSolver {
  std::vector<StageFactor> datas;
  Kkt0 kkt0;
  Mat thGrad, thHess;
  const Problem *problem;
  explicit ProximalRiccatiSolver(const Problem &p);
  bool backward(double mudyn, double mueq);
  bool forward(std::vector<Mat> &xs, std::vector<Mat> &us, std::vector<Mat> &vs,
               std::vector<Mat> &lbdas, const Mat *theta = nullptr) const;
};

/// dense-riccati.hpp:28-58 / dense-riccati.hxx, DenseKernel dense-kernel.hpp:56-212 (oracle/dense.cpp).
/// Shares StageFactor::{ff, fb, fth} (block rows [K; Z; L; Y]) and Value {Pmat = Pxx, pvec = px, Vxt = Pxt,
/// Vtt = Ptt, vt = pt} with the proximal solver.  Reads Gv.
struct RiccatiSolverDense {
  std::vector<StageFactor> datas;
  Kkt0 kkt0;
  Mat thGrad, thHess;
  const Problem *problem;
  explicit RiccatiSolverDense(const Problem &p);
  bool backward(double mudyn, double mueq);
  bool forward(std::vector<Mat> &xs, std::vector<Mat> &us, std::vector<Mat> &vs,
               std::vector<Mat> &lbdas, const Mat *theta = nullptr) const;
};

/// work.hpp:17-22
inline std::array<int, 2> get_work(int horz, int tid, int nthreads) {
  return {(int)((long)tid * (horz + 1) / nthreads), (int)((long)(tid + 1) * (horz + 1) / nthreads)};
}

/// block-tridiagonal.hpp:80-137 / :50-75 / :144-179
bool symmetricBlockTridiagSolve(std::vector<Mat> &sub, std::vector<Mat> &diag,
                                const std::vector<Mat> &super, std::vector<Mat> &rhs,
                                std::vector<BunchKaufman> &facs);
void blockTridiagMatMul(const std::vector<Mat> &sub, const std::vector<Mat> &diag,
                        const std::vector<Mat> &super, const std::vector<Mat> &b,
                        std::vector<Mat> &c, double beta);
void blockTridiagRefinementStep(const std::vector<Mat> &upFacs, const std::vector<Mat> &super,
                                const std::vector<BunchKaufman> &facs, std::vector<Mat> &rhs);

/// parallel-solver.hpp:24-108 / parallel-solver.hxx. MUTATES the problem like the
/// reference (adds the nth = nx parameterization to every leg but the last).
struct ParallelRiccatiSolver {
  std::vector<StageFactor> datas;
  int numThreads;
  std::vector<Mat> sub, diag, super;     // condensedKktSystem
  std::vector<Mat> diagFacs, upFacs;     // condensedFacs
  std::vector<BunchKaufman> ldlt;
  std::vector<Mat> rhs, sol, err;        // condensedKktRhs / Solution / Err
  double condensedThreshold = 1e-11;
  int refinementStepsTaken = 0;
  bool tridiagOk = true;
  Problem *problem;
  ParallelRiccatiSolver(Problem &p, int num_threads);
  static void setupKnot(Knot &knot, double mudyn);   // parallel-solver.hpp:43-51
  void assembleCondensedSystem(double mudyn);        // parallel-solver.hxx:63-109
  bool backward(double mudyn, double mueq);          // :111-179
  bool forward(std::vector<Mat> &xs, std::vector<Mat> &us, std::vector<Mat> &vs,
               std::vector<Mat> &lbdas) const;       // :181-215
  void collapseFeedback();                           // parallel-solver.hpp:55-66
};

/// utils.hpp:139-168
void lqrInitializeSolution(const Problem &p, std::vector<Mat> &xs, std::vector<Mat> &us,
                           std::vector<Mat> &vs, std::vector<Mat> &lbdas);
/// utils.hxx:94-192 -> {dyn, cstr, dual}
std::array<double, 3> lqrComputeKktError(const Problem &p, const std::vector<Mat> &xs,
                                         const std::vector<Mat> &us, const std::vector<Mat> &vs,
                                         const std::vector<Mat> &lbdas, double mudyn,
                                         double mueq, const Mat *theta);
/// utils.hpp:71-84
int lqrNumRows(const Problem &p);
/// utils.hxx:194-267
bool lqrDenseMatrix(const Problem &p, double mudyn, double mueq, Mat &mat, Mat &rhs);

} // namespace gar_oracle
