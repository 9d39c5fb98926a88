// kernels.cuh — shared declarations between the CUDA translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/garb200.h"

namespace garb200 {

/// work.hpp:17-22 — balanced leg partition [beg, end) of the T+1 knots.
__host__ __device__ inline void get_work(int horz, int tid, int nthreads, int &beg, int &end) {
  beg = (int)((long long)tid * (horz + 1) / nthreads);
  end = (int)((long long)(tid + 1) * (horz + 1) / nthreads);
}

/// Diagnostic tuning knobs (garb200_set_tuning): read by the LAUNCHERS only, fixed per solver.
struct Tune {
  int group = 0;          // problems per CTA of the nx = 36 sweep (0/1: one CTA per problem, default; >= 2: 3 in lockstep)
  int einv_mode = 0;      // 0: default pre-pass, 1: register Gauss-Jordan, 2: LU in shared memory (nx = 76)
  int einv_minb = 6;      // min CTAs/SM hint of the small-matrix pre-pass
  int fwd_stages = 2;     // TMA ring depth of the forward rollout
  int condensed = 0;      // 0: default, 1: Bunch-Kaufman kernel only, 2: specialised kernel only
  int c3_eg = 1;          // (36,12,6,4): E^{-1} read from L2 instead of staged
  int legs_generic = 0;   // leg mode on the generic kernel
  int explicit_dyn = 1;   // use the E = -I fast path when every knot qualifies
};

/// Everything a kernel needs to know about one solver (passed by value).
struct DevParams {
  int nx, nu, nc, nth;  // layout dims (nth: user parameter dim; 0 in leg mode)
  int hleg;             // parameter dim of the non-last legs (nx) in leg mode, else 0
  int T, nc0, batch, num_legs;
  int shift;            // ring offset of the time axis (cycleAppend)
  int store_all_values;
  int explicit_dyn;     // 1: every E_t == -I exactly (E and E^{-1} are never read by the specialised sweeps)
  int uniform_knots;    // 1: every knot t < T has (nu, nc), the terminal knot has nu = nc = 0
  int nslots;           // value slots per problem (T+1 when storing all, else #legs or 1)
  const int *nu_t, *nc_t;         // [T+1] logical-time dims
  double *knots;                  // [batch][T+1][KL.stride]   (physical time order)
  double *G0, *g0;                // [batch][nc0*nx], [batch][nc0]
  double *gains;                  // [batch][T+1][GL.stride]   (physical time order)
  double *einv;                   // [batch][T][nx*nx] E^{-1} of every knot (fast path pre-pass) or nullptr
  double *values;                 // [batch][nslots][VL.stride]
  double *kkt0_ff, *kkt0_fth, *thGrad, *thHess;  // per problem
  double *xs, *us, *vs, *lbdas;   // solver-owned trajectories
  double *theta;                  // [batch][nth] device copy (forward)
  double *gv;                     // dense solver only: Gv [batch][T+1][nc*nth] (col-major nc x nth, logical time order) or nullptr
  int dense;                      // 1: RiccatiSolverDense (dense.cu) instead of the proximal sweep
  // condensed system of the parallel-in-time solver (leg mode), per problem:
  // (blocks themselves are read in place from value / knot records; only factors are stored)
  double *cdiagF, *csubF, *crhs, *csol, *cerr, *csubd;
  int *cpiv;
  int32_t *flags, *refine_steps;
  long long *phase_cycles;        // [16] accumulated phase timers of the specialised sweep, or nullptr
  double mudyn, mueq;
  Tune tune;
  void *side_stream = nullptr, *side_e0 = nullptr, *side_e1 = nullptr;   // leg mode (host side only)
  garb200_knot_layout KL;
  garb200_gain_layout GL;   // built with nth_eff = (leg mode ? nx : nth)
  garb200_value_layout VL;  // same
};

__host__ __device__ inline int phys_t(const DevParams &p, int t) {
  // logical knot t -> physical record index; the terminal knot never moves
  return (t >= p.T || p.T == 0) ? t : (t + p.shift) % p.T;
}
__host__ __device__ inline int value_slot(const DevParams &p, int t) {
  if (p.store_all_values) return t;
  if (p.num_legs < 2) return t == 0 ? 0 : -1;
  for (int j = 0; j < p.num_legs; j++) {
    int b, e; get_work(p.T, j, p.num_legs, b, e);
    if (b == t) return j;
  }
  return -1;
}

struct LaunchCfg { int threads; size_t smem; int grid; };

// launchers (each returns the number of kernels it launched, < 0 on error)
int launch_backward(const DevParams &p, cudaStream_t st);
// dimension-specialised sweep (sweep_fast.cu): -3 = no specialisation for these dims
int launch_backward_fast(const DevParams &p, const int *h_nu_t, const int *h_nc_t, cudaStream_t st);
bool fast_path_available(int nx, int nu, int nc, int nth, int num_legs);
// streaming variant for records larger than shared memory (sweep_big.cu): -3 = no instantiation
int launch_backward_big(const DevParams &p, cudaStream_t st);
bool big_path_available(int nx, int nu, int nc, int nth);
int launch_forward(const DevParams &p, const double *d_theta, double *xs, double *us, double *vs,
                   double *lbdas, cudaStream_t st);
int launch_condensed(const DevParams &p, cudaStream_t st);
int launch_collapse(const DevParams &p, cudaStream_t st);
int launch_kkt_error(const DevParams &p, const double *d_theta, double mudyn, double mueq,
                     double *d_out, cudaStream_t st);
int launch_generate(const DevParams &p, uint64_t seed, int dist, cudaStream_t st);
// *d_flag = 1 iff E_t == -I exactly for every problem and knot t < T (one pass over the E blocks)
int launch_detect_explicit(const DevParams &p, int *d_flag, cudaStream_t st);
int fp64_peak(int kind, double *tflops);
// dense.cu: RiccatiSolverDense (one BK of the (nu + nc + 2 nx)^2 stage KKT matrix per knot)
int launch_backward_dense(const DevParams &p, cudaStream_t st);
// pad.cu: exact embedding of a shape without an instantiated sweep into the next instantiated one
bool padded_shape_for(int nx, int nu, int nc, int nth, int out[4]);
int launch_pad_problem(const DevParams &u, const DevParams &p, cudaStream_t st);
int launch_unpad_gains(const DevParams &u, const DevParams &p, cudaStream_t st);
int launch_unpad_traj(const DevParams &u, const DevParams &p, double *xs, double *us, double *vs, double *lbdas, cudaStream_t st);
int launch_pad_theta(const double *th, double *thp, int batch, int h, int H, cudaStream_t st);
size_t dense_smem_bytes(const DevParams &p);
// consumer.cu: device-side producer / consumers around the solve
int launch_pack_from_device(const DevParams &p, const garb200_device_problem_view &v, double preg, cudaStream_t st);
int launch_linear_step(const DevParams &p, double alpha, const double *const cur[4], double *const out[4], cudaStream_t st);
int launch_rollout(const DevParams &p, double alpha, const double *dxs, double *dus, double *dvs, double *dlams, cudaStream_t st);

} // namespace garb200
