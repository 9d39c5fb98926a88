/* garb200.h — C ABI of the B200-native batched backend for aligator's `gar` LQ solver.
 *
 * Drop-in boundary: `gar::RiccatiSolverBase<double>`
 * (/root/reference/include/aligator/gar/riccati-base.hpp:12-34) and its concrete
 * solvers ProximalRiccatiSolver (proximal-riccati.hpp:12-45) and ParallelRiccatiSolver
 * (parallel-solver.hpp:24-108).  Every entry point below names the reference interface
 * it replaces.  Plain pointers and sizes only: no Eigen, no torch, no C++ types.
 *
 * One solver handle owns `batch` INDEPENDENT LQ problems of identical dimensions
 * (the reference solves one problem per solver object).  All device and pinned memory
 * is allocated in garb200_create(); nothing is allocated afterwards — the reference's
 * ALIGATOR_NOMALLOC_SCOPED contract (proximal-riccati.hxx:35, parallel-solver.hxx:115).
 *
 * Error convention: every call returns a status (0 = GARB200_OK).  The reference's
 * `bool backward()` becomes a status plus a per-problem int32 flag array
 * (garb200_get_flags).  There is NO CPU fallback: without a CUDA device every compute
 * call returns GARB200_ERR_CUDA.
 *
 * Record formats: include/garb200_layout.h.  Trajectory arrays, per problem:
 *   xs    (T+1) x nx                         xs[t]  at t*nx
 *   us    (T+1) x nu   (rows of knots with nu_t < nu are zero padded; us[T] unused when
 *                       the terminal knot has nu = 0, utils.hpp:163-165)
 *   vs    (T+1) x nc
 *   lbdas nc0 + T*nx   lbdas[0] (nc0 entries, utils.hpp:150) then lbdas[t] at nc0+(t-1)*nx
 * and problem b follows problem b-1 contiguously in every array.
 */
#ifndef GARB200_H
#define GARB200_H

#include <stddef.h>
#include <stdint.h>
#include "garb200_layout.h"

#ifdef __cplusplus
extern "C" {
#endif

#define GARB200_OK 0
#define GARB200_ERR_INVALID 1   /* bad argument / unsupported dimensions            */
#define GARB200_ERR_CUDA 2      /* CUDA runtime error (see garb200_last_error)       */
#define GARB200_ERR_NUMERIC 3   /* >= 1 problem hit a zero pivot / non-finite value  */
#define GARB200_ERR_STATE 4     /* call order violated (e.g. forward before backward)*/

/* per-problem flag bits (garb200_get_flags) */
#define GARB200_FLAG_LU_SINGULAR 1   /* PartialPivLU(E) met a zero pivot (riccati-kernel.hxx:172) */
#define GARB200_FLAG_LLT_NOT_PD 2    /* LLT(I + mu*Ptilde) met a non-positive pivot (:185)       */
#define GARB200_FLAG_BK_SINGULAR 4   /* BunchKaufman -> NumericalIssue (bunchkaufman.hpp:56-57)  */
#define GARB200_FLAG_NONFINITE 8     /* NaN/Inf in the produced gains                            */
#define GARB200_FLAG_TRIDIAG_FAIL 16 /* block-tridiagonal BK failed (block-tridiagonal.hpp:102)  */

/* garb200_config.flags */
#define GARB200_STORE_ALL_VALUES 1   /* keep StageFactor::vm of EVERY knot (default: knot 0 and
                                        leg heads only — what the algorithm itself consumes)  */

#define GARB200_FORCE_GENERIC 2      /* never use the dimension-specialised kernels (testing)  */
#define GARB200_EXPLICIT_DYNAMICS 8  /* the caller PROMISES E_t == -I at every knot (explicit dynamics,
                                        explicit-dynamics.hxx:32): E is never read and the device-side
                                        check that garb200_backward otherwise runs after every change
                                        of the problem is skipped                                  */
#define GARB200_DENSE_KERNEL 16      /* RiccatiSolverDense (dense-riccati.hpp:28-58): one Bunch-Kaufman
                                        factorization of the (nu + nc + 2 nx)^2 stage KKT matrix per knot
                                        (dense-kernel.hpp:56-212) instead of the proximal sweep; reads Gv
                                        (garb200_upload_gv); the slow, sturdy path; num_legs must be 0   */
#define GARB200_NO_PADDING 32        /* do not embed a shape without an instantiated sweep into the next
                                        instantiated one (testing; see garb200_padded_shape)            */
#define GARB200_PHASE_TIMERS 4       /* diagnostic: per-phase clock64 timers in the specialised
                                        sweep (garb200_phase_cycles); slower instrumented kernel */

typedef struct garb200_solver garb200_solver; /* opaque */

typedef struct garb200_config {
  int nx;          /* state dim (nx2 == nx for every knot)                              */
  int nu;          /* max control dim over knots                                        */
  int nc;          /* max equality-constraint dim over knots                            */
  int nth;         /* parameter dim (LqrProblemTpl::ntheta, lqr-problem.hpp:250)        */
  int horizon;     /* T: the problem has T+1 knots (LqrProblemTpl::horizon)             */
  int nc0;         /* rows of the initial constraint G0 x0 + g0 = 0 (nc0())             */
  int batch;       /* number of independent problems                                    */
  int num_legs;    /* 0/1: ProximalRiccatiSolver; >= 2: ParallelRiccatiSolver with that
                      many parallel-in-time legs (`num_threads`, parallel-solver.hpp:38) */
  int device;      /* CUDA device ordinal                                               */
  int flags;       /* GARB200_STORE_ALL_VALUES | ...                                    */
  const int *nu_t; /* optional [T+1] per-knot nu (NULL: nu for t < T, nu_term at T)     */
  const int *nc_t; /* optional [T+1] per-knot nc (NULL: nc for t < T, nc_term at T)     */
  int nu_term;     /* used when nu_t == NULL (generated problems: 0, test_util.cpp:60)  */
  int nc_term;     /* used when nc_t == NULL                                            */
} garb200_config;

/* One knot as the reference stores it: 16 column-major matrices, each with its own
 * pointer (`ManagedMatrix::data()`, memory/managed-matrix.hpp:279-282) — so a binding can
 * pass LqrKnotTpl fields without copying.  NULL = all zeros. Gv is not read
 * (riccati-kernel.hxx:128,272). */
typedef struct garb200_knot_view {
  int nu, nc; /* this knot's own dims (<= config) */
  const double *Q, *S, *R, *q, *r;
  const double *A, *B, *E, *f;
  const double *C, *D, *d;
  const double *Gth, *Gx, *Gu, *gamma;
} garb200_knot_view;

/* ---- life cycle -------------------------------------------------------------------
 * replaces ProximalRiccatiSolver(const LqrProblemTpl&)  proximal-riccati.hxx:12-30
 *      and ParallelRiccatiSolver(LqrProblemTpl&, uint)  parallel-solver.hxx:18-49      */
int garb200_create(const garb200_config *cfg, garb200_solver **out);
void garb200_destroy(garb200_solver *s);
const char *garb200_last_error(void);
const char *garb200_version(void);

/* Diagnostic tuning knobs, read by the kernel launchers of this solver only (no environment
 * variables anywhere in the library).  Names: "group" (problems per CTA of the nx = 36 sweep:
 * 0/1 one CTA per problem (default), >= 2 three problems in lockstep), "einv_mode", "einv_minb", "fwd_stages", "condensed",
 * "c3_eg", "legs_generic", "explicit_dyn" (0: never take the E = -I path). */
int garb200_set_tuning(garb200_solver *s, const char *name, int value);
/* Run subsequent kernels/copies on this CUDA stream (a cudaStream_t, passed as void*). */
int garb200_set_stream(garb200_solver *s, void *cuda_stream);
/* Block until everything queued by this solver has finished. */
int garb200_synchronize(garb200_solver *s);

/* layouts actually used by this solver (gains/values use nth_eff = nx in leg mode) */
int garb200_get_layouts(const garb200_solver *s, garb200_knot_layout *knot,
                        garb200_gain_layout *gain, garb200_value_layout *value);
/* same, without a solver (pure function of the dims) */
void garb200_layouts(int nx, int nu, int nc, int nth, int *knot17, int *gain5, int *value8);

/* ---- problem data (the solver keeps a device-resident copy; the reference re-reads
 * `problem_->stages` on every backward(), proximal-riccati.hxx:37)  -------------------- */
/* whole batch, already packed: knots [batch][T+1][stride], G0 [batch][nc0*nx] col-major,
 * g0 [batch][nc0]; host pointers */
int garb200_upload_packed(garb200_solver *s, const double *knots, const double *G0,
                          const double *g0);
/* one knot from reference-style matrices into the pinned staging buffer */
int garb200_stage_knot(garb200_solver *s, int problem, int t, const garb200_knot_view *k);
int garb200_stage_initial(garb200_solver *s, int problem, const double *G0, const double *g0);
/* H2D of everything staged so far */
int garb200_commit(garb200_solver *s);
/* Gv of every knot for the DENSE solver (GARB200_DENSE_KERNEL; the proximal kernel never reads it,
 * riccati-kernel.hxx:128,272 vs dense-kernel.hpp:76,134): host [batch][T+1][nc * nth], column-major nc x nth
 * blocks with the solver's max nc as leading dimension, logical time order; NULL = zeros */
int garb200_upload_gv(garb200_solver *s, const double *gv);
/* raw device pointers of the resident problem (for upstream device producers) */
int garb200_device_problem(garb200_solver *s, double **d_knots, double **d_G0, double **d_g0);
/* ---- device-resident producer ------------------------------------------------
 * SolverProxDDPTpl::updateLQSubproblem          solvers/proxddp/solver-proxddp.hxx:802-873
 * A batched caller keeps its derivatives on the device; this packs them into the resident problem with
 * one kernel (no host staging, no H2D): knot.A = Jx, knot.B = Ju, knot.E = Jy, knot.f = Lds[t+1],
 * knot.Q = Lxx (+ preg on the diagonal), knot.S = Lxu, knot.R = Luu (+ preg), q, r, C, D, d (:824-850),
 * G0 = init Jx, g0 = Lds[0] (:867-868).  Element (i, j) of knot t of problem b of a field F is read at
 * F.ptr[b * stride_b + t * stride_t + i + j * ld] (column-major blocks, any strides); F.ptr == NULL means
 * zeros — except E, where NULL means EXPLICIT dynamics, E = -I (explicit-dynamics.hxx:32).  Blocks of a
 * knot are read with that knot's own (nu_t, nc_t). */
typedef struct garb200_device_field {
  const double *ptr;
  long long stride_b, stride_t;
  int ld;
} garb200_device_field;
typedef struct garb200_device_problem_view {
  garb200_device_field Q, S, R, q, r, A, B, E, f, C, D, d, Gth, Gx, Gu, gamma, G0, g0;
} garb200_device_problem_view;
int garb200_pack_from_device(garb200_solver *s, const garb200_device_problem_view *v, double preg);

/* fill the resident problem with synthetic random data ON DEVICE (bench data source;
 * distributions of tests/gar/test_util.cpp:13-66; dist 0 = "ref", 1 = "explicit" E = -I with the
 * generator's zeroed Q block (nonconvex stages: Bunch-Kaufman at every knot), 2 = "explicit_psd" E = -I
 * with convex stage costs) */
int garb200_generate(garb200_solver *s, uint64_t seed, int dist);
/* D2H of the packed problem of `count` problems starting at `first` */
int garb200_download_packed(garb200_solver *s, int first, int count, double *knots, double *G0,
                            double *g0);

/* ---- the hot path -------------------------------------------------------------------
 * bool RiccatiSolverBase::backward(mudyn, mueq)          riccati-base.hpp:18
 *   ProximalRiccatiSolver::backward                      proximal-riccati.hxx:32-64
 *   ParallelRiccatiSolver::backward                      parallel-solver.hxx:111-179   */
int garb200_backward(garb200_solver *s, double mudyn, double mueq);

/* bool RiccatiSolverBase::forward(xs, us, vs, lbdas, theta)   riccati-base.hpp:20-23
 * theta: host, [batch][nth] or NULL (ignored in leg mode, parallel-solver.hxx:184).
 * Outputs: HOST pointers (any may be NULL to skip the D2H copy of that array).         */
int garb200_forward(garb200_solver *s, const double *theta, double *xs, double *us,
                    double *vs, double *lbdas);
/* same with DEVICE pointers for theta and outputs (NULL output = solver-owned buffer) */
int garb200_forward_device(garb200_solver *s, const double *d_theta, double *d_xs,
                           double *d_us, double *d_vs, double *d_lbdas);
/* solver-owned device trajectory buffers (valid after forward*) */
int garb200_device_trajectory(garb200_solver *s, double **d_xs, double **d_us, double **d_vs,
                              double **d_lbdas);

/* solver-owned device buffer for theta, [batch][nth] (input of garb200_forward_device) */
int garb200_device_theta(garb200_solver *s, double **d_theta);

/* ---- several GPUs of one box behind ONE logical solver (SURVEY 8(b) "multi-GPU by opts.devices[]", 8(e)) ----
 * The reference has no counterpart (its ParallelRiccatiSolver parallelises ONE problem over CPU threads);
 * independent problems shard with no communication inside backward / forward.  cfg->batch is the GLOBAL
 * batch, cfg->device is ignored; GPU g of devices[] owns problems [g B / G, (g + 1) B / G).  One host
 * thread drives all devices (every call is asynchronous).  The only exchange is the gather of the
 * trajectories in garb200_multi_forward: root_device < 0 -> HOST output pointers for the global batch (each
 * device copies its slice to its offset); root_device >= 0 -> DEVICE pointers on that GPU, slices arrive by
 * cudaMemcpyPeerAsync (copy engines over NVLink: no kernel competes with the next sweep for SMs).
 * Parallel-in-time mode does not shard (one problem = one GPU): create one solver per device instead. */
typedef struct garb200_multi garb200_multi;
int garb200_multi_create(const garb200_config *cfg, const int *devices, int ndev, garb200_multi **out);
void garb200_multi_destroy(garb200_multi *m);
int garb200_multi_num_shards(const garb200_multi *m);
int garb200_multi_shard(garb200_multi *m, int i, garb200_solver **solver, int *device, int *first, int *count);
int garb200_multi_upload_packed(garb200_multi *m, const double *knots, const double *G0, const double *g0);
int garb200_multi_generate(garb200_multi *m, uint64_t seed, int dist);
int garb200_multi_backward(garb200_multi *m, double mudyn, double mueq);
int garb200_multi_forward(garb200_multi *m, const double *theta, int root_device, double *xs, double *us,
                          double *vs, double *lbdas);
int garb200_multi_synchronize(garb200_multi *m);
int garb200_multi_kkt_error(garb200_multi *m, double mudyn, double mueq, const double *theta, double *out);

/* ---- device-resident consumers -----------------------------------------------
 * SolverProxDDPTpl::tryLinearStep (vector spaces)    solver-proxddp.hxx:104-148
 *   trial_* = cur_* + alpha * (direction of the last forward*, solver-owned) for xs, us, vs, lbdas in one
 *   pass; DEVICE pointers in the trajectory layouts at the top of this file; a NULL cur/trial pair is skipped. */
int garb200_linear_step(garb200_solver *s, double alpha, const double *d_xs, const double *d_us,
                        const double *d_vs, const double *d_lbdas, double *d_trial_xs,
                        double *d_trial_us, double *d_trial_vs, double *d_trial_lbdas);
/* the linear part of SolverProxDDPTpl::tryNonlinearRollout   solver-proxddp.hxx:372-391, 431-446
 *   dus[t] = alpha kff_t + K_t dxs[t], dvs[t] = alpha zff_t + Z_t dxs[t], dlbdas[t+1] = alpha lff_t + L_t dxs[t]
 * for every knot from the gains records in place (getFeedforward / getFeedback never leave the device);
 * d_dxs [batch][T+1][nx] is the caller's state deviation; dlbdas[0] (nc0 entries) is not written. */
int garb200_rollout_device(garb200_solver *s, double alpha, const double *d_dxs, double *d_dus,
                           double *d_dvs, double *d_dlbdas);

/* ---- solver state readers ------------------------------------------------------------
 * getFeedforward(i) / getFeedback(i)                      riccati-base.hpp:30-31
 * datas[t].{ff,fb,fth}                                    riccati-kernel.hpp:76-78      */
int garb200_get_gains(garb200_solver *s, int problem, int t, double *ff, double *fb_rowmajor,
                      double *fth_rowmajor);
/* whole packed gains record(s) [count][T+1][gain stride] */
int garb200_download_gains(garb200_solver *s, int first, int count, double *gains);
int garb200_device_gains(garb200_solver *s, double **d_gains);
/* datas[t].vm (riccati-kernel.hpp:39-60); one packed value record. Needs
 * GARB200_STORE_ALL_VALUES unless t is 0 (or a leg head in leg mode). */
int garb200_get_value(garb200_solver *s, int problem, int t, double *value_record);
/* kkt0.ff [nx+nc0], kkt0.fth [(nx+nc0) x nth] row-major, thGrad [nth], thHess [nth x nth]
 * column-major (proximal-riccati.hpp:39-41) */
int garb200_get_initial(garb200_solver *s, int problem, double *kkt0_ff, double *kkt0_fth,
                        double *thGrad, double *thHess);
int garb200_get_flags(garb200_solver *s, int32_t *flags /* [batch] */);
/* leg mode: number of iterative-refinement steps taken per problem (parallel-solver.hxx:155-175) */
int garb200_get_refinement_steps(garb200_solver *s, int32_t *steps /* [batch] */);

/* void collapseFeedback()                                 parallel-solver.hpp:55-66     */
int garb200_collapse_feedback(garb200_solver *s);
/* void cycleAppend(const LqrKnot&)   proximal-riccati.hxx:80-88, parallel-solver.hxx:217-238
 * Rotates the time axis of problem AND factor storage left by one (ring offset, no copy)
 * and installs `knots[b]` (one view per problem, or one shared view if `shared` != 0)
 * as the new knot T-1.                                                                  */
int garb200_cycle_append(garb200_solver *s, const garb200_knot_view *knots, int shared);

/* ---- metric: lqrComputeKktError                         utils.hxx:94-192
 * On the trajectories of the last forward*; out [batch][3] = {dyn, cstr, dual} (host). */
int garb200_kkt_error(garb200_solver *s, double mudyn, double mueq, const double *theta,
                      double *out);

/* ---- allocation-free variant of the metric -------------------------------------------------
 * This entry point has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernel and same numbers as garb200_kkt_error, but no
 * solver handle and no allocation of any kind: the kernel's scratch (per-knot nu_t / nc_t arrays,
 * [batch][3] result staging — cudaMalloc'd by garb200_create on the handle path) lives in ONE static
 * __device__ array of GARB200_NOALLOC_SCRATCH_BYTES per GPU, held from library load until exit.
 *   dims      nx, nu, nc, nth, horizon, nc0, batch, device, nu_t/nc_t (host, or NULL -> nu_term/
 *             nc_term) are read; num_legs and flags are ignored
 *   d_*       DEVICE pointers: packed knots [batch][T+1][stride] in logical time order
 *             (garb200_layout.h), G0 [batch][nc0*nx], g0 [batch][nc0], trajectories laid out as at the
 *             top of this file, theta [batch][nth] or NULL
 *   out       HOST [batch][3] = {dyn, cstr, dual}; the call returns after the stream has drained
 * Limit: 256-rounded 8 (T+1) bytes + 24 batch bytes <= GARB200_NOALLOC_SCRATCH_BYTES, otherwise
 * GARB200_ERR_INVALID with the sizes in garb200_last_error() (e.g. T = 50: batch <= 43 667).
 * The array is shared by every call on a device and keeps the last call's contents (each call
 * rewrites what it reads and zeroes the result staging first): calls that OVERLAP on one GPU race. */
#define GARB200_NOALLOC_SCRATCH_BYTES (1u << 20)
int garb200_kkt_error_noalloc(const garb200_config *dims, const double *d_knots, const double *d_G0,
                              const double *d_g0, const double *d_xs, const double *d_us,
                              const double *d_vs, const double *d_lbdas, const double *d_theta,
                              double mudyn, double mueq, void *cuda_stream, double *out);

/* Host trajectory -> solver-owned device buffers (so garb200_kkt_error can score a solution
 * that was produced elsewhere, like the reference's free function does). */
int garb200_upload_trajectory(garb200_solver *s, const double *xs, const double *us,
                              const double *vs, const double *lbdas);

/* ---- introspection for bench / roofline ---------------------------------------------- */
/* number of kernels launched by this solver since creation */
int64_t garb200_kernel_launches(const garb200_solver *s);
/* device-event time (ms) of the last backward and forward kernels sequences */
int garb200_last_timings(garb200_solver *s, float *backward_ms, float *forward_ms);
/* 1 if the last backward ran the dimension-specialised (DMMA) sweep, 0 for the generic kernel */
int garb200_used_fast_path(const garb200_solver *s);
/* Shape generality: a problem whose (nx, nu, nc, nth) — or whose ragged per-knot nu_t / nc_t — has no
 * instantiated sweep is solved by the specialised sweep of the next instantiated shape through EXACT
 * zero / identity padding (csrc/pad.cu); layouts seen by the caller stay the caller's.  Returns 1 and the
 * internal shape {nx, nu, nc, nth} if this solver is padded, else 0. */
int garb200_padded_shape(const garb200_solver *s, int shape4[4]);
/* 1 if it ran with the explicit-dynamics path (every E_t == -I: no E^-1 pre-pass, E never read) */
int garb200_used_explicit_path(const garb200_solver *s);
/* Diagnostic: accumulated clock64() cycles of thread 0 of every CTA per phase of the specialised
 * sweep since the last call (solver created with GARB200_PHASE_TIMERS in config.flags). */
int garb200_phase_cycles(garb200_solver *s, long long *out16);
/* Per-call device timing log (bench / roofline): after _start, every backward and forward
 * brackets its kernels with CUDA events on the solver stream (up to `capacity` of each);
 * _read synchronises, returns the per-call durations in ms and stops logging. */
int garb200_timing_log_start(garb200_solver *s, int capacity);
int garb200_timing_log_read(garb200_solver *s, float *backward_ms, float *forward_ms, int *count);
/* FP64 micro-benchmarks used as roofline denominators: returns TFLOP/s
 * kind 0 = DFMA register loop, 1 = DMMA mma.sync.m8n8k4.f64 loop */
int garb200_fp64_peak(int device, int kind, double *tflops);

#ifdef __cplusplus
}
#endif
#endif /* GARB200_H */
