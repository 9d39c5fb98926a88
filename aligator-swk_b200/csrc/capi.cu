// capi.cu — the C ABI of include/garb200.h: host-side solver object over the CUDA kernels.
// No CPU fallback lives here: every compute entry point launches sm_100a kernels or
// fails with GARB200_ERR_CUDA.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: ranges show up in nsys / ncu, cost nothing without a tool attached
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "kernels.cuh"

namespace garb200 { size_t backward_smem_bytes(const DevParams &p); }
using namespace garb200;

// NVTX ranges around the phases of a solve (the reference's Tracy zones: riccati-kernel.hxx:67,93,156,170,257,308)
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
namespace garb200 { int set_error(int code, const std::string &msg) { return fail(code, msg); } }   // for noalloc.cu
#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(GARB200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));      \
  } while (0)

struct garb200_solver {
  garb200_config cfg{};
  DevParams p{};
  std::vector<int> nu_t, nc_t;
  std::vector<void *> dev_allocs;
  cudaStream_t stream = nullptr;
  double *h_stage = nullptr;     // pinned staging of the packed problem (small batches only)
  double *h_G0 = nullptr, *h_g0 = nullptr;
  double *h_scratch = nullptr;   // pinned: one gains/value record, flags, kkt out ...
  size_t scratch_doubles = 0;
  double *h_cycle = nullptr;     // pinned [batch][kstride] for cycleAppend
  double *d_kkt_out = nullptr;
  int64_t launches = 0;
  bool has_problem = false, did_backward = false, did_forward = false, used_fast = false;
  cudaEvent_t ev[4]{};
  std::vector<cudaEvent_t> tlog;   // timing log: 4 events per logged solve (bwd begin/end, fwd begin/end)
  int tlog_cap = 0, tlog_bwd = 0, tlog_fwd = 0;
  size_t knots_doubles = 0;
  // staging bookkeeping: only knots staged since the last commit are uploaded (a knot installed
  // another way — upload_packed, generate, device producers, cycle_append — is never overwritten)
  std::vector<uint8_t> dirty;    // [batch][T+1] physical knot slots staged since the last commit
  std::vector<uint8_t> dirty0;   // [batch] initial condition staged since the last commit
  bool any_dirty = false;
  cudaEvent_t ev_h2d = nullptr;  // recorded after every async H2D that reads a pinned staging buffer
  bool h2d_pending = false;
  double *h_gains = nullptr;     // pinned mirror of the gains records (getFeedback/getFeedforward cache)
  bool gains_fresh = false;
  bool problem_dirty = true;     // the resident problem changed since E == -I was last checked
  // shape generality (pad.cu): this solver keeps the USER-shaped problem / gains / trajectories and owns an
  // inner solver of the covering instantiated shape that does the work
  garb200_solver *inner = nullptr;
  bool inner_dirty = true;       // the user-shaped problem changed since it was last padded into the inner one
  bool user_gains_fresh = false; // p.gains (user layout) mirrors the inner gains of the last factorization
  int pad_shape[4] = {0, 0, 0, 0};
  int *d_expl = nullptr;         // device flag written by detect_explicit_kernel
  cudaStream_t side = nullptr;   // leg mode: the last leg runs beside the coupled legs
  cudaEvent_t side_e0 = nullptr, side_e1 = nullptr;
  int nblk = 0;                  // blocks of the condensed system (2 * legs)
  size_t csys_len = 0;           // nc0 + (nblk-1) nx
};

static const size_t kMaxPinnedStage = (size_t)1 << 30;  // 1 GiB
static const size_t kMaxPinnedGains = (size_t)256 << 20; // 256 MiB

// The pinned staging buffers may still be the source of an in-flight H2D copy (commit / cycle_append
// are asynchronous): wait for it before the host overwrites them.
static int wait_staging(garb200_solver *s) {
  if (s->h2d_pending) {
    if (cudaEventSynchronize(s->ev_h2d) != cudaSuccess) { g_err = "cudaEventSynchronize (staging) failed"; return GARB200_ERR_CUDA; }
    s->h2d_pending = false;
  }
  return GARB200_OK;
}

template <class T> static int dmalloc(garb200_solver *s, T **ptr, size_t count) {
  *ptr = nullptr;
  if (count == 0) count = 1;
  void *q = nullptr;
  CK(cudaMalloc(&q, count * sizeof(T)));
  s->dev_allocs.push_back(q);
  *ptr = static_cast<T *>(q);
  return GARB200_OK;
}
#define TRY(x) do { int rc_ = (x); if (rc_ != GARB200_OK) return rc_; } while (0)
// padded shape: the user-layout gains records are produced from the inner solver's on first use after a
// factorization (one kernel), then every reader below works on them unchanged
static int sync_user_gains(garb200_solver *s) {
  if (!s->inner || s->user_gains_fresh) return GARB200_OK;
  const int rc = launch_unpad_gains(s->p, s->inner->p, s->stream);
  if (rc < 0) return fail(GARB200_ERR_CUDA, "unpad_gains launch failed");
  s->launches += rc;
  s->user_gains_fresh = true;
  return GARB200_OK;
}

extern "C" {

const char *garb200_last_error(void) { return g_err.c_str(); }
const char *garb200_version(void) { return "garb200 0.1 (sm_100a)"; }

void garb200_layouts(int nx, int nu, int nc, int nth, int *knot, int *gain, int *value) {
  const garb200_knot_layout L = garb200_make_knot_layout(nx, nu, nc, nth);
  const int k[17] = {L.Q, L.S, L.R, L.q, L.r, L.A, L.B, L.E, L.f, L.C, L.D, L.d,
                     L.Gth, L.Gx, L.Gu, L.gamma, L.stride};
  for (int i = 0; i < 17; i++) knot[i] = k[i];
  const garb200_gain_layout G = garb200_make_gain_layout(nx, nu, nc, nth);
  gain[0] = G.nrows; gain[1] = G.ff; gain[2] = G.fb; gain[3] = G.fth; gain[4] = G.stride;
  const garb200_value_layout V = garb200_make_value_layout(nx, nth);
  const int v[8] = {V.Pmat, V.pvec, V.Vxx, V.vx, V.Vxt, V.Vtt, V.vt, V.stride};
  for (int i = 0; i < 8; i++) value[i] = v[i];
}

void garb200_destroy(garb200_solver *s) {
  if (!s) return;
  if (s->inner) { garb200_destroy(s->inner); s->inner = nullptr; }
  cudaSetDevice(s->cfg.device);
  if (s->stream) cudaStreamSynchronize(s->stream); else cudaDeviceSynchronize();
  for (void *q : s->dev_allocs) cudaFree(q);
  if (s->h_stage) cudaFreeHost(s->h_stage);
  if (s->h_G0) cudaFreeHost(s->h_G0);
  if (s->h_g0) cudaFreeHost(s->h_g0);
  if (s->h_scratch) cudaFreeHost(s->h_scratch);
  if (s->h_cycle) cudaFreeHost(s->h_cycle);
  if (s->h_gains) cudaFreeHost(s->h_gains);
  if (s->ev_h2d) cudaEventDestroy(s->ev_h2d);
  if (s->side) cudaStreamDestroy(s->side);
  if (s->side_e0) cudaEventDestroy(s->side_e0);
  if (s->side_e1) cudaEventDestroy(s->side_e1);
  for (auto &e : s->ev) if (e) cudaEventDestroy(e);
  for (auto &e : s->tlog) if (e) cudaEventDestroy(e);
  delete s;
}

int garb200_create(const garb200_config *cfg, garb200_solver **out) {
  if (!cfg || !out) return fail(GARB200_ERR_INVALID, "null argument");
  *out = nullptr;
  const garb200_config &c = *cfg;
  if (c.nx < 1 || c.nu < 0 || c.nc < 0 || c.nth < 0 || c.horizon < 0 || c.batch < 1 || c.nc0 < 0)
    return fail(GARB200_ERR_INVALID, "negative or zero dimension");
  if (c.nc0 > c.nx) return fail(GARB200_ERR_INVALID, "nc0 > nx is not supported");
  const bool legs = c.num_legs >= 2;
  if (legs && c.nth != 0)   // the reference overwrites the parameterization (parallel-solver.hxx:51-61)
    return fail(GARB200_ERR_INVALID, "parallel-in-time mode requires nth == 0");
  if (legs && c.num_legs > c.horizon + 1)
    return fail(GARB200_ERR_INVALID, "more legs than knots");
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (c.device < 0 || c.device >= ndev) return fail(GARB200_ERR_CUDA, "no such CUDA device");
  CK(cudaSetDevice(c.device));

  garb200_solver *s = new garb200_solver();
  s->cfg = c;
  const int T = c.horizon;
  s->nu_t.resize(T + 1); s->nc_t.resize(T + 1);
  for (int t = 0; t <= T; t++) {
    s->nu_t[t] = c.nu_t ? c.nu_t[t] : (t < T ? c.nu : c.nu_term);
    s->nc_t[t] = c.nc_t ? c.nc_t[t] : (t < T ? c.nc : c.nc_term);
    if (s->nu_t[t] < 0 || s->nu_t[t] > c.nu || s->nc_t[t] < 0 || s->nc_t[t] > c.nc) {
      delete s;
      return fail(GARB200_ERR_INVALID, "per-knot nu/nc outside [0, max]");
    }
  }
  s->cfg.nu_t = nullptr; s->cfg.nc_t = nullptr;
  DevParams &p = s->p;
  p.nx = c.nx; p.nu = c.nu; p.nc = c.nc; p.nth = c.nth;
  p.hleg = legs ? c.nx : 0;
  p.T = T; p.nc0 = c.nc0; p.batch = c.batch; p.num_legs = legs ? c.num_legs : 0;
  p.shift = 0;
  p.store_all_values = (c.flags & GARB200_STORE_ALL_VALUES) ? 1 : 0;
  p.explicit_dyn = 0;
  p.uniform_knots = (s->nu_t[T] == 0 && s->nc_t[T] == 0) ? 1 : 0;
  for (int t = 0; t < T; t++)
    if (s->nu_t[t] != c.nu || s->nc_t[t] != c.nc) p.uniform_knots = 0;
  p.nslots = p.store_all_values ? T + 1 : (legs ? c.num_legs : 1);
  const int nth_eff = legs ? c.nx : c.nth;
  p.KL = garb200_make_knot_layout(c.nx, c.nu, c.nc, c.nth);
  p.GL = garb200_make_gain_layout(c.nx, c.nu, c.nc, nth_eff);
  p.VL = garb200_make_value_layout(c.nx, nth_eff);
  p.mudyn = p.mueq = 0.0;
  p.dense = (c.flags & GARB200_DENSE_KERNEL) ? 1 : 0;
  p.gv = nullptr;
  if (p.dense && legs) { delete s; return fail(GARB200_ERR_INVALID, "GARB200_DENSE_KERNEL: the dense solver has no parallel-in-time mode"); }
  if (p.dense && dense_smem_bytes(p) > 227u * 1024u) {
    char buf[256];
    snprintf(buf, sizeof buf, "GARB200_DENSE_KERNEL: the (nu + nc + 2 nx)^2 stage KKT matrix needs %zu bytes of shared memory (> 227 KB)", dense_smem_bytes(p));
    delete s;
    return fail(GARB200_ERR_INVALID, buf);
  }
  // Shape generality (pad.cu): no instantiated sweep for these dims, or ragged per-knot dims -> the work is
  // done by an inner solver of the covering instantiated shape; this one keeps the user-shaped buffers.
  bool padded = false;
  if (!p.dense && !legs && !(c.flags & (GARB200_FORCE_GENERIC | GARB200_NO_PADDING)) && T >= 1 &&
      s->nu_t[T] == 0 && s->nc_t[T] == 0 &&
      !(p.uniform_knots && fast_path_available(c.nx, c.nu, c.nc, c.nth, 0)))
    padded = padded_shape_for(c.nx, c.nu, c.nc, c.nth, s->pad_shape) &&
             fast_path_available(s->pad_shape[0], s->pad_shape[1], s->pad_shape[2], s->pad_shape[3], 0);
  const bool want_fast = !padded && !p.dense && !(c.flags & GARB200_FORCE_GENERIC) && fast_path_available(c.nx, c.nu, c.nc, c.nth, c.num_legs);
  // (a shape with a specialised sweep may exceed the generic kernel's shared-memory plan: it then has
  //  no generic fallback and ragged knots are refused at backward())
  // A shape that neither a specialised sweep (native or by padding) nor the generic proximal kernel's
  // shared-memory plan can hold still has the dense per-knot KKT solver (csrc/dense.cu: same interface, same
  // records, same solution to round-off; slow but sturdy) as long as ITS matrix fits: taken automatically.
  if (!want_fast && !padded && !p.dense && !legs && !(c.flags & GARB200_FORCE_GENERIC) &&
      backward_smem_bytes(p) > 227u * 1024u && dense_smem_bytes(p) <= 227u * 1024u)
    p.dense = 1;
  if (!want_fast && !padded && !p.dense && backward_smem_bytes(p) > 227u * 1024u) {
    char buf[256];
    snprintf(buf, sizeof buf, "dimensions need %zu bytes of shared memory per CTA (> 227 KB): "
             "nx=%d nu=%d nc=%d nth=%d not supported by this kernel generation",
             backward_smem_bytes(p), c.nx, c.nu, c.nc, nth_eff);
    delete s;
    return fail(GARB200_ERR_INVALID, buf);
  }
  auto bail = [&](int rc) { garb200_destroy(s); return rc; };
#define TRYB(x) do { int rc_ = (x); if (rc_ != GARB200_OK) return bail(rc_); } while (0)
  const size_t B = c.batch, n = c.nx;
  s->knots_doubles = B * (T + 1) * p.KL.stride;
  int *d_nu = nullptr, *d_nc = nullptr;
  TRYB(dmalloc(s, &d_nu, T + 1)); TRYB(dmalloc(s, &d_nc, T + 1));
  if (cudaMemcpy(d_nu, s->nu_t.data(), (T + 1) * sizeof(int), cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(d_nc, s->nc_t.data(), (T + 1) * sizeof(int), cudaMemcpyHostToDevice) != cudaSuccess)
    return bail(fail(GARB200_ERR_CUDA, "memcpy of per-knot dims failed"));
  p.nu_t = d_nu; p.nc_t = d_nc;
  TRYB(dmalloc(s, &p.knots, s->knots_doubles));
  TRYB(dmalloc(s, &p.G0, B * c.nc0 * n));
  TRYB(dmalloc(s, &p.g0, B * c.nc0));
  TRYB(dmalloc(s, &p.gains, B * (T + 1) * p.GL.stride));
  TRYB(dmalloc(s, &p.values, B * p.nslots * p.VL.stride));
  p.einv = nullptr;
  if (want_fast)
    TRYB(dmalloc(s, &p.einv, B * (size_t)T * n * n));
  TRYB(dmalloc(s, &p.kkt0_ff, B * (n + c.nc0)));
  TRYB(dmalloc(s, &p.kkt0_fth, B * (n + c.nc0) * c.nth));
  TRYB(dmalloc(s, &p.thGrad, B * c.nth));
  TRYB(dmalloc(s, &p.thHess, B * c.nth * c.nth));
  TRYB(dmalloc(s, &p.xs, B * (T + 1) * n));
  TRYB(dmalloc(s, &p.us, B * (T + 1) * c.nu));
  TRYB(dmalloc(s, &p.vs, B * (T + 1) * c.nc));
  TRYB(dmalloc(s, &p.lbdas, B * (c.nc0 + (size_t)T * n)));
  TRYB(dmalloc(s, &p.theta, B * c.nth));
  if (p.dense && c.nc * c.nth > 0) {
    TRYB(dmalloc(s, &p.gv, B * (T + 1) * c.nc * c.nth));
    cudaMemset(p.gv, 0, B * (T + 1) * c.nc * c.nth * sizeof(double));
  }
  TRYB(dmalloc(s, &p.flags, B));
  TRYB(dmalloc(s, &p.refine_steps, B));
  TRYB(dmalloc(s, &s->d_kkt_out, B * 3));
  TRYB(dmalloc(s, &s->d_expl, 1));
  p.phase_cycles = nullptr;
  if (c.flags & GARB200_PHASE_TIMERS) { TRYB(dmalloc(s, &p.phase_cycles, 16)); cudaMemset(p.phase_cycles, 0, 16 * sizeof(long long)); }
  cudaMemset(p.flags, 0, B * sizeof(int32_t));
  cudaMemset(p.refine_steps, 0, B * sizeof(int32_t));
  cudaMemset(p.values, 0, B * p.nslots * p.VL.stride * sizeof(double));
  cudaMemset(p.gains, 0, B * (T + 1) * p.GL.stride * sizeof(double));
  if (legs) {
    s->nblk = 2 * c.num_legs;
    s->csys_len = c.nc0 + (size_t)(s->nblk - 1) * n;
    const size_t nb = s->nblk, bb = n * n;
    TRYB(dmalloc(s, &p.cdiagF, B * nb * bb));
    TRYB(dmalloc(s, &p.csubF, B * nb * bb));
    TRYB(dmalloc(s, &p.csubd, B * nb * n));
    TRYB(dmalloc(s, &p.cpiv, B * nb * n));
    TRYB(dmalloc(s, &p.crhs, B * s->csys_len));
    TRYB(dmalloc(s, &p.csol, B * s->csys_len));
    TRYB(dmalloc(s, &p.cerr, B * s->csys_len));
  }
  // pinned host buffers
  if (s->knots_doubles * sizeof(double) <= kMaxPinnedStage) {
    if (cudaMallocHost(&s->h_stage, std::max<size_t>(s->knots_doubles, 1) * sizeof(double)) != cudaSuccess ||
        cudaMallocHost(&s->h_G0, std::max<size_t>(B * c.nc0 * n, 1) * sizeof(double)) != cudaSuccess ||
        cudaMallocHost(&s->h_g0, std::max<size_t>(B * c.nc0, 1) * sizeof(double)) != cudaSuccess)
      return bail(fail(GARB200_ERR_CUDA, "cudaMallocHost (problem staging) failed"));
    memset(s->h_stage, 0, s->knots_doubles * sizeof(double));
    memset(s->h_G0, 0, B * c.nc0 * n * sizeof(double));
    memset(s->h_g0, 0, B * c.nc0 * sizeof(double));
  }
  s->scratch_doubles = std::max<size_t>({(size_t)p.GL.stride, (size_t)p.VL.stride, 3 * B, B,
                                         (size_t)(n + c.nc0) * (1 + c.nth) + c.nth * (1 + c.nth)}) + 16;
  if (cudaMallocHost(&s->h_scratch, s->scratch_doubles * sizeof(double)) != cudaSuccess)
    return bail(fail(GARB200_ERR_CUDA, "cudaMallocHost (scratch) failed"));
  if (B * p.KL.stride * sizeof(double) <= kMaxPinnedStage / 4)
    if (cudaMallocHost(&s->h_cycle, B * p.KL.stride * sizeof(double)) != cudaSuccess)
      return bail(fail(GARB200_ERR_CUDA, "cudaMallocHost (cycle) failed"));
  for (auto &e : s->ev)
    if (cudaEventCreate(&e) != cudaSuccess) return bail(fail(GARB200_ERR_CUDA, "cudaEventCreate failed"));
  if (cudaEventCreateWithFlags(&s->ev_h2d, cudaEventDisableTiming) != cudaSuccess)
    return bail(fail(GARB200_ERR_CUDA, "cudaEventCreate failed"));
  if (s->h_stage) { s->dirty.assign(B * (T + 1), 0); s->dirty0.assign(B, 0); }
  {
    const size_t gbytes = B * (T + 1) * p.GL.stride * sizeof(double);
    if (gbytes <= kMaxPinnedGains && cudaMallocHost(&s->h_gains, gbytes) != cudaSuccess)
      return bail(fail(GARB200_ERR_CUDA, "cudaMallocHost (gains mirror) failed"));
  }
  if (legs) {   // side stream of the last leg: created here, never lazily (no allocation after create)
    if (cudaStreamCreateWithFlags(&s->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->side_e0, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->side_e1, cudaEventDisableTiming) != cudaSuccess)
      return bail(fail(GARB200_ERR_CUDA, "side stream creation failed"));
    p.side_stream = s->side; p.side_e0 = s->side_e0; p.side_e1 = s->side_e1;
  }
  if (padded) {
    garb200_config ic = c;
    ic.nx = s->pad_shape[0]; ic.nu = s->pad_shape[1]; ic.nc = s->pad_shape[2]; ic.nth = s->pad_shape[3];
    ic.nc0 = c.nc0 + (ic.nx - c.nx);          // the padded states are pinned to zero by G0' = [G0 0; 0 I]
    ic.nu_t = nullptr; ic.nc_t = nullptr; ic.nu_term = 0; ic.nc_term = 0;
    ic.flags = c.flags | GARB200_NO_PADDING;
    TRYB(garb200_create(&ic, &s->inner));
    // numerical flags and refinement counters are per problem: read the inner solver's
    p.flags = s->inner->p.flags; p.refine_steps = s->inner->p.refine_steps;
  }
#undef TRYB
  *out = s;
  return GARB200_OK;
}

int garb200_padded_shape(const garb200_solver *s, int shape4[4]) {
  if (!s || !s->inner) return 0;
  if (shape4) for (int i = 0; i < 4; i++) shape4[i] = s->pad_shape[i];
  return 1;
}

int garb200_set_tuning(garb200_solver *s, const char *name, int value) {
  if (!s || !name) return fail(GARB200_ERR_INVALID, "null argument");
  if (s->inner) TRY(garb200_set_tuning(s->inner, name, value));
  Tune &t = s->p.tune;
  const std::string k(name);
  if (k == "group") t.group = value;
  else if (k == "einv_mode") t.einv_mode = value;
  else if (k == "einv_minb") t.einv_minb = value;
  else if (k == "fwd_stages") t.fwd_stages = value;
  else if (k == "condensed") t.condensed = value;
  else if (k == "c3_eg") t.c3_eg = value;
  else if (k == "legs_generic") t.legs_generic = value;
  else if (k == "explicit_dyn") { t.explicit_dyn = value; s->problem_dirty = s->inner_dirty = true; }
  else return fail(GARB200_ERR_INVALID, "unknown tuning knob: " + k);
  return GARB200_OK;
}

int garb200_set_stream(garb200_solver *s, void *cuda_stream) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  s->stream = static_cast<cudaStream_t>(cuda_stream);
  if (s->inner) s->inner->stream = s->stream;
  return GARB200_OK;
}

static int check_flags(garb200_solver *s) {
  // caller has synchronised the stream
  const size_t B = s->cfg.batch;
  int32_t *hf = reinterpret_cast<int32_t *>(s->h_scratch);
  CK(cudaMemcpyAsync(hf, s->p.flags, B * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  int bad = 0;
  for (size_t i = 0; i < B; i++) bad += hf[i] != 0;
  if (bad) {
    char buf[128];
    snprintf(buf, sizeof buf, "%d of %zu problems raised a numerical flag (garb200_get_flags)", bad, B);
    return fail(GARB200_ERR_NUMERIC, buf);
  }
  return GARB200_OK;
}

int garb200_synchronize(garb200_solver *s) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  CK(cudaSetDevice(s->cfg.device));
  CK(cudaStreamSynchronize(s->stream));
  return s->did_backward ? check_flags(s) : GARB200_OK;
}

int garb200_get_layouts(const garb200_solver *s, garb200_knot_layout *knot,
                        garb200_gain_layout *gain, garb200_value_layout *value) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (knot) *knot = s->p.KL;
  if (gain) *gain = s->p.GL;
  if (value) *value = s->p.VL;
  return GARB200_OK;
}

// ---- problem data ----------------------------------------------------------------------
// Packed knots between the caller's LOGICAL time order and the solver's physical (ring-
// offset, see garb200_cycle_append) order: logical [0, T-shift) <-> physical [shift, T),
// logical [T-shift, T) <-> physical [0, shift), the terminal knot never moves.
static int copy_knots_rotated(garb200_solver *s, double *dev, double *host, int first, int count,
                              cudaMemcpyKind kind) {
  const DevParams &p = s->p;
  const size_t ks = p.KL.stride, T = p.T, sh = p.shift, pitch = (T + 1) * ks * sizeof(double);
  double *d0 = dev + (size_t)first * (T + 1) * ks;
  auto seg = [&](size_t logical, size_t phys, size_t nk) -> cudaError_t {
    if (nk == 0 || count == 0) return cudaSuccess;
    double *d = d0 + phys * ks, *h = host + logical * ks;
    return kind == cudaMemcpyHostToDevice
               ? cudaMemcpy2DAsync(d, pitch, h, pitch, nk * ks * sizeof(double), count, kind, s->stream)
               : cudaMemcpy2DAsync(h, pitch, d, pitch, nk * ks * sizeof(double), count, kind, s->stream);
  };
  if (sh == 0) { CK(seg(0, 0, T + 1)); return GARB200_OK; }
  CK(seg(0, sh, T - sh));
  CK(seg(T - sh, 0, sh));
  CK(seg(T, T, 1));
  return GARB200_OK;
}
int garb200_upload_packed(garb200_solver *s, const double *knots, const double *G0,
                          const double *g0) {
  NvtxRange nvtx_("garb200::upload_packed (H2D)");
  if (!s || !knots) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(s->cfg.device));
  const DevParams &p = s->p;
  const size_t B = p.batch, n = p.nx;
  TRY(copy_knots_rotated(s, p.knots, const_cast<double *>(knots), 0, p.batch, cudaMemcpyHostToDevice));
  if (G0) CK(cudaMemcpyAsync(p.G0, G0, B * p.nc0 * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  if (g0) CK(cudaMemcpyAsync(p.g0, g0, B * p.nc0 * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  s->has_problem = true;
  s->problem_dirty = s->inner_dirty = true;
  return GARB200_OK;
}

static void pack_block(double *dst, const double *src, int count) {
  if (count <= 0) return;
  if (src) memcpy(dst, src, (size_t)count * sizeof(double));
  else memset(dst, 0, (size_t)count * sizeof(double));
}

static void pack_knot(const garb200_knot_layout &L, const garb200_knot_view &k, double *rec) {
  const int n = L.nx, m = k.nu, c = k.nc, h = L.nth;
  pack_block(rec + L.Q, k.Q, n * n); pack_block(rec + L.S, k.S, n * m); pack_block(rec + L.R, k.R, m * m);
  pack_block(rec + L.q, k.q, n);     pack_block(rec + L.r, k.r, m);
  pack_block(rec + L.A, k.A, n * n); pack_block(rec + L.B, k.B, n * m); pack_block(rec + L.E, k.E, n * n);
  pack_block(rec + L.f, k.f, n);
  pack_block(rec + L.C, k.C, c * n); pack_block(rec + L.D, k.D, c * m); pack_block(rec + L.d, k.d, c);
  pack_block(rec + L.Gth, k.Gth, h * h); pack_block(rec + L.Gx, k.Gx, n * h);
  pack_block(rec + L.Gu, k.Gu, m * h);   pack_block(rec + L.gamma, k.gamma, h);
}

int garb200_stage_knot(garb200_solver *s, int problem, int t, const garb200_knot_view *k) {
  NvtxRange nvtx_("garb200::stage_knot (host pack)");
  if (!s || !k) return fail(GARB200_ERR_INVALID, "null argument");
  const DevParams &p = s->p;
  if (!s->h_stage) return fail(GARB200_ERR_STATE, "batch too large for pinned staging: use garb200_upload_packed");
  if (problem < 0 || problem >= p.batch || t < 0 || t > p.T) return fail(GARB200_ERR_INVALID, "index out of range");
  if (k->nu != s->nu_t[t] || k->nc != s->nc_t[t]) return fail(GARB200_ERR_INVALID, "knot dims differ from the solver's");
  TRY(wait_staging(s));
  const size_t slot = (size_t)problem * (p.T + 1) + phys_t(p, t);
  pack_knot(p.KL, *k, s->h_stage + slot * p.KL.stride);
  s->dirty[slot] = 1; s->any_dirty = true;
  return GARB200_OK;
}

int garb200_stage_initial(garb200_solver *s, int problem, const double *G0, const double *g0) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  const DevParams &p = s->p;
  if (!s->h_stage) return fail(GARB200_ERR_STATE, "batch too large for pinned staging");
  if (problem < 0 || problem >= p.batch) return fail(GARB200_ERR_INVALID, "index out of range");
  TRY(wait_staging(s));
  pack_block(s->h_G0 + (size_t)problem * p.nc0 * p.nx, G0, p.nc0 * p.nx);
  pack_block(s->h_g0 + (size_t)problem * p.nc0, g0, p.nc0);
  s->dirty0[problem] = 1; s->any_dirty = true;
  return GARB200_OK;
}

int garb200_commit(garb200_solver *s) {
  NvtxRange nvtx_("garb200::commit (H2D of staged knots)");
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (!s->h_stage) return fail(GARB200_ERR_STATE, "batch too large for pinned staging");
  CK(cudaSetDevice(s->cfg.device));
  const DevParams &p = s->p;
  const size_t B = p.batch, n = p.nx, ks = p.KL.stride, total = B * (p.T + 1);
  if (s->any_dirty) {
    // contiguous runs of staged knot slots -> one copy each (a fully staged batch is ONE copy)
    for (size_t i = 0; i < total;) {
      if (!s->dirty[i]) { i++; continue; }
      size_t j = i;
      while (j < total && s->dirty[j]) s->dirty[j++] = 0;
      CK(cudaMemcpyAsync(p.knots + i * ks, s->h_stage + i * ks, (j - i) * ks * sizeof(double), cudaMemcpyHostToDevice, s->stream));
      i = j;
    }
    for (size_t i = 0; i < B;) {
      if (!s->dirty0[i]) { i++; continue; }
      size_t j = i;
      while (j < B && s->dirty0[j]) s->dirty0[j++] = 0;
      if (p.nc0 > 0) {
        CK(cudaMemcpyAsync(p.G0 + i * p.nc0 * n, s->h_G0 + i * p.nc0 * n, (j - i) * p.nc0 * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
        CK(cudaMemcpyAsync(p.g0 + i * p.nc0, s->h_g0 + i * p.nc0, (j - i) * p.nc0 * sizeof(double), cudaMemcpyHostToDevice, s->stream));
      }
      i = j;
    }
    CK(cudaEventRecord(s->ev_h2d, s->stream));
    s->h2d_pending = true;
    s->any_dirty = false;
    s->problem_dirty = s->inner_dirty = true;
  }
  s->has_problem = true;
  return GARB200_OK;
}

int garb200_upload_gv(garb200_solver *s, const double *gv) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  const DevParams &p = s->p;
  if (!p.dense) return fail(GARB200_ERR_STATE, "Gv is only read by the dense solver: create with GARB200_DENSE_KERNEL");
  const size_t len = (size_t)p.batch * (p.T + 1) * p.nc * p.nth;
  if (len == 0) return GARB200_OK;
  CK(cudaSetDevice(s->cfg.device));
  if (gv) CK(cudaMemcpyAsync(p.gv, gv, len * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  else CK(cudaMemsetAsync(p.gv, 0, len * sizeof(double), s->stream));
  CK(cudaStreamSynchronize(s->stream));     // gv may be pageable caller memory
  return GARB200_OK;
}

int garb200_device_problem(garb200_solver *s, double **d_knots, double **d_G0, double **d_g0) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (d_knots) *d_knots = s->p.knots;
  if (d_G0) *d_G0 = s->p.G0;
  if (d_g0) *d_g0 = s->p.g0;
  s->has_problem = true;  // the caller fills it on device
  s->problem_dirty = s->inner_dirty = true;
  return GARB200_OK;
}

int garb200_pack_from_device(garb200_solver *s, const garb200_device_problem_view *v, double preg) {
  NvtxRange nvtx_("garb200::pack_from_device");
  if (!s || !v) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(s->cfg.device));
  const int rc = launch_pack_from_device(s->p, *v, preg, s->stream);
  if (rc < 0) return fail(GARB200_ERR_CUDA, "pack_from_device launch failed");
  s->launches += rc;
  s->has_problem = true;
  s->problem_dirty = s->inner_dirty = true;
  return GARB200_OK;
}

int garb200_linear_step(garb200_solver *s, double alpha, const double *d_xs, const double *d_us,
                        const double *d_vs, const double *d_lbdas, double *d_trial_xs, double *d_trial_us,
                        double *d_trial_vs, double *d_trial_lbdas) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (!s->did_forward) return fail(GARB200_ERR_STATE, "linear_step needs a forward() into the solver-owned trajectory");
  CK(cudaSetDevice(s->cfg.device));
  const double *cur[4] = {d_xs, d_us, d_vs, d_lbdas};
  double *out[4] = {d_trial_xs, d_trial_us, d_trial_vs, d_trial_lbdas};
  const int rc = launch_linear_step(s->p, alpha, cur, out, s->stream);
  if (rc < 0) return fail(GARB200_ERR_CUDA, "linear_step launch failed");
  s->launches += rc;
  return GARB200_OK;
}

int garb200_rollout_device(garb200_solver *s, double alpha, const double *d_dxs, double *d_dus, double *d_dvs,
                           double *d_dlbdas) {
  if (!s || !d_dxs || !d_dlbdas) return fail(GARB200_ERR_INVALID, "null argument");
  if ((s->p.nu > 0 && !d_dus) || (s->p.nc > 0 && !d_dvs)) return fail(GARB200_ERR_INVALID, "null output for a non-empty block");
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "no factorization yet");
  if (s->p.num_legs >= 2) return fail(GARB200_ERR_STATE, "the nonlinear rollout is not supported with the parallel-in-time solver (solver-proxddp.hxx:170-173)");
  CK(cudaSetDevice(s->cfg.device));
  TRY(sync_user_gains(s));
  const int rc = launch_rollout(s->p, alpha, d_dxs, d_dus, d_dvs, d_dlbdas, s->stream);
  if (rc < 0) return fail(rc == -1 ? GARB200_ERR_INVALID : GARB200_ERR_CUDA, "rollout launch failed");
  s->launches += rc;
  return GARB200_OK;
}

int garb200_generate(garb200_solver *s, uint64_t seed, int dist) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  CK(cudaSetDevice(s->cfg.device));
  const int rc = launch_generate(s->p, seed, dist, s->stream);
  if (rc < 0) return fail(rc == -1 ? GARB200_ERR_INVALID : GARB200_ERR_CUDA, "generate kernel launch failed");
  s->launches += rc;
  s->has_problem = true;
  s->problem_dirty = s->inner_dirty = true;
  return GARB200_OK;
}

int garb200_download_packed(garb200_solver *s, int first, int count, double *knots, double *G0,
                            double *g0) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  const DevParams &p = s->p;
  if (first < 0 || count < 0 || first + count > p.batch) return fail(GARB200_ERR_INVALID, "range");
  CK(cudaSetDevice(s->cfg.device));
  const size_t n = p.nx;
  if (knots) TRY(copy_knots_rotated(s, p.knots, knots, first, count, cudaMemcpyDeviceToHost));
  if (G0) CK(cudaMemcpyAsync(G0, p.G0 + (size_t)first * p.nc0 * n, (size_t)count * p.nc0 * n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (g0) CK(cudaMemcpyAsync(g0, p.g0 + (size_t)first * p.nc0, (size_t)count * p.nc0 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return GARB200_OK;
}

// ---- the hot path ----------------------------------------------------------------------
int garb200_backward(garb200_solver *s, double mudyn, double mueq) {
  NvtxRange nvtx_("garb200::backward (E^-1 pre-pass + sweep [+ condensed])");
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (!s->has_problem) return fail(GARB200_ERR_STATE, "no problem uploaded");
  CK(cudaSetDevice(s->cfg.device));
  DevParams &p = s->p;
  p.mudyn = mudyn; p.mueq = mueq;
  if (s->inner) {   // padded shape: re-embed the problem if it changed, then the inner solver's sweep
    garb200_solver *in = s->inner;
    if (mueq == 0.0)   // an empty padded constraint row has the pivot -mueq
      for (int t = 0; t < p.T; t++)
        if (s->nc_t[t] < in->p.nc)
          return fail(GARB200_ERR_INVALID, "mueq == 0 with constraint rows padded to the instantiated shape: use mueq > 0 or GARB200_NO_PADDING");
    CK(cudaEventRecord(s->ev[0], s->stream));
    const bool logp = s->tlog_bwd < s->tlog_cap;
    if (logp) CK(cudaEventRecord(s->tlog[4 * s->tlog_bwd + 0], s->stream));
    if (s->inner_dirty) {
      in->p.shift = p.shift;
      const int rcp = launch_pad_problem(p, in->p, s->stream);
      if (rcp < 0) return fail(GARB200_ERR_CUDA, "pad_problem launch failed");
      s->launches += rcp;
      in->has_problem = true; in->problem_dirty = true;
      s->inner_dirty = false;
    }
    const int64_t l0 = in->launches;
    TRY(garb200_backward(in, mudyn, mueq));
    s->launches += in->launches - l0;
    s->used_fast = in->used_fast;
    p.explicit_dyn = in->p.explicit_dyn;
    CK(cudaEventRecord(s->ev[1], s->stream));
    if (logp) { CK(cudaEventRecord(s->tlog[4 * s->tlog_bwd + 1], s->stream)); s->tlog_bwd++; }
    s->did_backward = true; s->did_forward = false; s->gains_fresh = false; s->user_gains_fresh = false;
    return GARB200_OK;
  }
  CK(cudaMemsetAsync(p.flags, 0, (size_t)p.batch * sizeof(int32_t), s->stream));
  // Explicit dynamics (E_t == -I for every knot, explicit-dynamics.hxx:32): checked on the device whenever
  // the resident problem changed (one pass over the E blocks + a 4-byte read back), or promised by the
  // caller with GARB200_EXPLICIT_DYNAMICS.  Only the specialised sweeps have such a path.
  if (s->problem_dirty) {
    p.explicit_dyn = 0;
    if (p.einv != nullptr && p.tune.explicit_dyn && p.T >= 1) {
      if (s->cfg.flags & GARB200_EXPLICIT_DYNAMICS) p.explicit_dyn = 1;
      else {
        const int rcd = launch_detect_explicit(p, s->d_expl, s->stream);
        if (rcd < 0) return fail(GARB200_ERR_CUDA, "explicit-dynamics detection launch failed");
        s->launches += rcd;
        int *hflag = reinterpret_cast<int *>(s->h_scratch);
        CK(cudaMemcpyAsync(hflag, s->d_expl, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        p.explicit_dyn = *hflag != 0;
      }
    }
    s->problem_dirty = false;
  }
  CK(cudaEventRecord(s->ev[0], s->stream));
  const bool logb = s->tlog_bwd < s->tlog_cap;
  if (logb) CK(cudaEventRecord(s->tlog[4 * s->tlog_bwd + 0], s->stream));
  int rc = p.dense ? launch_backward_dense(p, s->stream)
                   : (p.einv != nullptr) ? launch_backward_fast(p, s->nu_t.data(), s->nc_t.data(), s->stream) : -3;
  s->used_fast = (rc >= 0) && !p.dense;
  if (rc == -3) rc = launch_backward(p, s->stream);
  if (rc < 0) return fail(rc == -1 ? GARB200_ERR_INVALID : GARB200_ERR_CUDA,
                          std::string("backward launch failed: ") + cudaGetErrorString(cudaGetLastError()));
  s->launches += rc;
  if (p.num_legs >= 2) {
    rc = launch_condensed(p, s->stream);
    if (rc < 0) return fail(GARB200_ERR_CUDA, "condensed-system launch failed");
    s->launches += rc;
  }
  CK(cudaEventRecord(s->ev[1], s->stream));
  if (logb) { CK(cudaEventRecord(s->tlog[4 * s->tlog_bwd + 1], s->stream)); s->tlog_bwd++; }
  s->did_backward = true;
  s->did_forward = false;
  s->gains_fresh = false;
  return GARB200_OK;
}

int garb200_forward_device(garb200_solver *s, const double *d_theta, double *d_xs, double *d_us,
                           double *d_vs, double *d_lbdas) {
  NvtxRange nvtx_("garb200::forward (rollout kernel)");
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "forward before backward");
  CK(cudaSetDevice(s->cfg.device));
  const DevParams &p = s->p;
  CK(cudaEventRecord(s->ev[2], s->stream));
  const bool logf = s->tlog_fwd < s->tlog_cap;
  if (logf) CK(cudaEventRecord(s->tlog[4 * s->tlog_fwd + 2], s->stream));
  if (s->inner) {   // padded shape: theta -> internal width, inner rollout, trajectories back to the user's widths
    garb200_solver *in = s->inner;
    const double *th_in = nullptr;
    if (d_theta && p.nth > 0) {
      const int rt = launch_pad_theta(d_theta, in->p.theta, p.batch, p.nth, in->p.nth, s->stream);
      if (rt < 0) return fail(GARB200_ERR_CUDA, "pad_theta launch failed");
      s->launches += rt;
      th_in = in->p.theta;
    }
    const int64_t l0 = in->launches;
    TRY(garb200_forward_device(in, th_in, nullptr, nullptr, nullptr, nullptr));
    s->launches += in->launches - l0;
    const int ru = launch_unpad_traj(p, in->p, d_xs ? d_xs : p.xs, d_us ? d_us : p.us, d_vs ? d_vs : p.vs,
                                     d_lbdas ? d_lbdas : p.lbdas, s->stream);
    if (ru < 0) return fail(GARB200_ERR_CUDA, "unpad_traj launch failed");
    s->launches += ru;
    CK(cudaEventRecord(s->ev[3], s->stream));
    if (logf) { CK(cudaEventRecord(s->tlog[4 * s->tlog_fwd + 3], s->stream)); s->tlog_fwd++; }
    s->did_forward = (!d_xs && !d_us && !d_vs && !d_lbdas);
    return GARB200_OK;
  }
  const int rc = launch_forward(p, (p.nth > 0 && p.num_legs < 2) ? d_theta : nullptr,
                                d_xs ? d_xs : p.xs, d_us ? d_us : p.us, d_vs ? d_vs : p.vs,
                                d_lbdas ? d_lbdas : p.lbdas, s->stream);
  if (rc < 0) return fail(rc == -1 ? GARB200_ERR_INVALID : GARB200_ERR_CUDA, "forward launch failed");
  s->launches += rc;
  CK(cudaEventRecord(s->ev[3], s->stream));
  if (logf) { CK(cudaEventRecord(s->tlog[4 * s->tlog_fwd + 3], s->stream)); s->tlog_fwd++; }
  s->did_forward = (!d_xs && !d_us && !d_vs && !d_lbdas);
  return GARB200_OK;
}

int garb200_forward(garb200_solver *s, const double *theta, double *xs, double *us, double *vs,
                    double *lbdas) {
  NvtxRange nvtx_("garb200::forward + D2H");
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "forward before backward");
  CK(cudaSetDevice(s->cfg.device));
  const DevParams &p = s->p;
  const size_t B = p.batch, n = p.nx, T = p.T;
  const double *d_theta = nullptr;
  if (theta && p.nth > 0 && p.num_legs < 2) {
    CK(cudaMemcpyAsync(p.theta, theta, B * p.nth * sizeof(double), cudaMemcpyHostToDevice, s->stream));
    d_theta = p.theta;
  }
  TRY(garb200_forward_device(s, d_theta, nullptr, nullptr, nullptr, nullptr));
  if (xs) CK(cudaMemcpyAsync(xs, p.xs, B * (T + 1) * n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (us && p.nu) CK(cudaMemcpyAsync(us, p.us, B * (T + 1) * p.nu * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (vs && p.nc) CK(cudaMemcpyAsync(vs, p.vs, B * (T + 1) * p.nc * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  if (lbdas) CK(cudaMemcpyAsync(lbdas, p.lbdas, B * (p.nc0 + T * n) * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return check_flags(s);
}

int garb200_device_theta(garb200_solver *s, double **d_theta) {
  if (!s || !d_theta) return fail(GARB200_ERR_INVALID, "null argument");
  *d_theta = s->p.theta;
  return GARB200_OK;
}

int garb200_device_trajectory(garb200_solver *s, double **d_xs, double **d_us, double **d_vs,
                              double **d_lbdas) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (d_xs) *d_xs = s->p.xs;
  if (d_us) *d_us = s->p.us;
  if (d_vs) *d_vs = s->p.vs;
  if (d_lbdas) *d_lbdas = s->p.lbdas;
  return GARB200_OK;
}

// ---- solver state readers ----------------------------------------------------------------
static int fetch(garb200_solver *s, double *dst_host, const double *src_dev, size_t count) {
  CK(cudaMemcpyAsync(dst_host, src_dev, count * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return GARB200_OK;
}

int garb200_get_gains(garb200_solver *s, int problem, int t, double *ff, double *fb, double *fth) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  const DevParams &p = s->p;
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "no factorization yet");
  if (problem < 0 || problem >= p.batch || t < 0 || t > p.T) return fail(GARB200_ERR_INVALID, "index out of range");
  CK(cudaSetDevice(s->cfg.device));
  TRY(sync_user_gains(s));
  // updateGains() reads every knot of a problem right after backward (solver-proxddp.hxx:316-334): the
  // first read after a factorization mirrors ALL gains records in one bulk D2H, the others are memcpys
  const size_t slot = (size_t)problem * (p.T + 1) + phys_t(p, t);
  const double *rec_h;
  if (s->h_gains) {
    if (!s->gains_fresh) {
      TRY(fetch(s, s->h_gains, p.gains, (size_t)p.batch * (p.T + 1) * p.GL.stride));
      s->gains_fresh = true;
    }
    rec_h = s->h_gains + slot * p.GL.stride;
  } else {
    TRY(fetch(s, s->h_scratch, p.gains + slot * p.GL.stride, p.GL.stride));
    rec_h = s->h_scratch;
  }
  const int nr = s->nu_t[t] + s->nc_t[t] + 2 * p.nx;
  if (ff) memcpy(ff, rec_h + p.GL.ff, nr * sizeof(double));
  if (fb) memcpy(fb, rec_h + p.GL.fb, (size_t)nr * p.nx * sizeof(double));
  if (fth && p.GL.nth) memcpy(fth, rec_h + p.GL.fth, (size_t)nr * p.GL.nth * sizeof(double));
  return GARB200_OK;
}

int garb200_download_gains(garb200_solver *s, int first, int count, double *gains) {
  NvtxRange nvtx_("garb200::download_gains (D2H)");
  if (!s || !gains) return fail(GARB200_ERR_INVALID, "null argument");
  const DevParams &p = s->p;
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "no factorization yet");
  if (first < 0 || count < 0 || first + count > p.batch) return fail(GARB200_ERR_INVALID, "range");
  CK(cudaSetDevice(s->cfg.device));
  TRY(sync_user_gains(s));
  const size_t ps = (size_t)(p.T + 1) * p.GL.stride;
  if (p.shift == 0) return fetch(s, gains, p.gains + first * ps, count * ps);
  for (int b = first; b < first + count; b++)      // undo the ring offset: logical time order out
    for (int t = 0; t <= p.T; t++)
      CK(cudaMemcpyAsync(gains + ((size_t)(b - first) * (p.T + 1) + t) * p.GL.stride,
                         p.gains + ((size_t)b * (p.T + 1) + phys_t(p, t)) * p.GL.stride,
                         p.GL.stride * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return GARB200_OK;
}

int garb200_device_gains(garb200_solver *s, double **d_gains) {
  if (!s || !d_gains) return fail(GARB200_ERR_INVALID, "null argument");
  if (s->inner && s->did_backward) { CK(cudaSetDevice(s->cfg.device)); TRY(sync_user_gains(s)); }
  *d_gains = s->p.gains;
  return GARB200_OK;
}

int garb200_get_value(garb200_solver *s, int problem, int t, double *value_record) {
  if (!s || !value_record) return fail(GARB200_ERR_INVALID, "null argument");
  const DevParams &p = s->p;
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "no factorization yet");
  if (problem < 0 || problem >= p.batch || t < 0 || t > p.T) return fail(GARB200_ERR_INVALID, "index out of range");
  const int slot = value_slot(p, t);
  if (slot < 0) return fail(GARB200_ERR_STATE, "value function of this knot was not kept: create with GARB200_STORE_ALL_VALUES");
  CK(cudaSetDevice(s->cfg.device));
  if (s->inner) {   // padded shape: the real part of the inner value function, in the user's record layout
    const DevParams &q = s->inner->p;
    std::vector<double> rec(q.VL.stride);
    TRY(fetch(s, rec.data(), q.values + ((size_t)problem * q.nslots + slot) * q.VL.stride, q.VL.stride));
    const int n = p.nx, N = q.nx, h = p.VL.nth, H = q.VL.nth;
    const garb200_value_layout &U = p.VL, &I = q.VL;
    for (int j = 0; j < n; j++)
      for (int i = 0; i < n; i++) {
        value_record[U.Pmat + i + j * n] = rec[I.Pmat + i + j * N];
        value_record[U.Vxx + i + j * n] = rec[I.Vxx + i + j * N];
      }
    for (int i = 0; i < n; i++) { value_record[U.pvec + i] = rec[I.pvec + i]; value_record[U.vx + i] = rec[I.vx + i]; }
    for (int a = 0; a < h; a++) {
      for (int i = 0; i < n; i++) value_record[U.Vxt + i + a * n] = rec[I.Vxt + i + a * N];
      for (int b2 = 0; b2 < h; b2++) value_record[U.Vtt + b2 + a * h] = rec[I.Vtt + b2 + a * H];
      value_record[U.vt + a] = rec[I.vt + a];
    }
    return GARB200_OK;
  }
  return fetch(s, value_record, p.values + ((size_t)problem * p.nslots + slot) * p.VL.stride, p.VL.stride);
}

int garb200_get_initial(garb200_solver *s, int problem, double *kkt0_ff, double *kkt0_fth,
                        double *thGrad, double *thHess) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  const DevParams &p = s->p;
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "no factorization yet");
  if (p.num_legs >= 2) return fail(GARB200_ERR_STATE, "the parallel-in-time solver has no kkt0 (parallel-solver.hpp)");
  if (problem < 0 || problem >= p.batch) return fail(GARB200_ERR_INVALID, "index out of range");
  CK(cudaSetDevice(s->cfg.device));
  const size_t s0 = p.nx + p.nc0, h = p.nth, b = problem;
  if (s->inner) {   // padded shape: rows [x0(n) | pad | lambda0(c0) | pad] x (H columns) -> the user's
    const DevParams &q = s->inner->p;
    const size_t N = q.nx, S0 = q.nx + q.nc0, H = q.nth, n = p.nx, c0 = p.nc0;
    std::vector<double> ff(S0), fth(S0 * H + 1), tg(H + 1), thh(H * H + 1);
    TRY(fetch(s, ff.data(), q.kkt0_ff + b * S0, S0));
    if (h) {
      TRY(fetch(s, fth.data(), q.kkt0_fth + b * S0 * H, S0 * H));
      TRY(fetch(s, tg.data(), q.thGrad + b * H, H));
      TRY(fetch(s, thh.data(), q.thHess + b * H * H, H * H));
    }
    auto irow = [&](size_t r) { return r < n ? r : N + (r - n); };
    for (size_t r = 0; r < n + c0; r++) {
      if (kkt0_ff) kkt0_ff[r] = ff[irow(r)];
      if (kkt0_fth) for (size_t a = 0; a < h; a++) kkt0_fth[r * h + a] = fth[irow(r) * H + a];
    }
    for (size_t a = 0; a < h; a++) {
      if (thGrad) thGrad[a] = tg[a];
      if (thHess) for (size_t c2 = 0; c2 < h; c2++) thHess[c2 + a * h] = thh[c2 + a * H];
    }
    return GARB200_OK;
  }
  if (kkt0_ff) TRY(fetch(s, kkt0_ff, p.kkt0_ff + b * s0, s0));
  if (kkt0_fth && h) TRY(fetch(s, kkt0_fth, p.kkt0_fth + b * s0 * h, s0 * h));
  if (thGrad && h) TRY(fetch(s, thGrad, p.thGrad + b * h, h));
  if (thHess && h) TRY(fetch(s, thHess, p.thHess + b * h * h, h * h));
  return GARB200_OK;
}

int garb200_get_flags(garb200_solver *s, int32_t *flags) {
  if (!s || !flags) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(s->cfg.device));
  CK(cudaMemcpyAsync(flags, s->p.flags, (size_t)s->p.batch * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return GARB200_OK;
}

int garb200_get_refinement_steps(garb200_solver *s, int32_t *steps) {
  if (!s || !steps) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(s->cfg.device));
  CK(cudaMemcpyAsync(steps, s->p.refine_steps, (size_t)s->p.batch * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return GARB200_OK;
}

int garb200_collapse_feedback(garb200_solver *s) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  if (s->p.num_legs < 2) return GARB200_OK;   // RiccatiSolverBase::collapseFeedback() {} (riccati-base.hpp:29)
  if (!s->did_backward) return fail(GARB200_ERR_STATE, "no factorization yet");
  CK(cudaSetDevice(s->cfg.device));
  const int rc = launch_collapse(s->p, s->stream);
  if (rc < 0) return fail(GARB200_ERR_CUDA, "collapse launch failed");
  s->launches += rc;
  s->gains_fresh = false;
  return GARB200_OK;
}

int garb200_cycle_append(garb200_solver *s, const garb200_knot_view *knots, int shared) {
  if (!s || !knots) return fail(GARB200_ERR_INVALID, "null argument");
  DevParams &p = s->p;
  if (p.T < 1) return fail(GARB200_ERR_INVALID, "cycle_append needs horizon >= 1");
  if (!s->h_cycle) return fail(GARB200_ERR_STATE, "batch too large for cycle_append staging");
  CK(cudaSetDevice(s->cfg.device));
  // the rotating knots must share dims (ring offset instead of the reference's copy-rotate)
  for (int t = 1; t < p.T; t++)
    if (s->nu_t[t] != s->nu_t[0] || s->nc_t[t] != s->nc_t[0])
      return fail(GARB200_ERR_INVALID, "cycle_append needs uniform nu/nc over knots 0..T-1");
  for (int b = 0; b < p.batch; b++) {
    const garb200_knot_view &k = shared ? knots[0] : knots[b];
    if (k.nu != s->nu_t[0] || k.nc != s->nc_t[0]) return fail(GARB200_ERR_INVALID, "knot dims differ from the solver's");
  }
  CK(cudaStreamSynchronize(s->stream));         // nothing in flight still reads the old offset or h_cycle
  s->h2d_pending = false;
  p.shift = (p.shift + 1) % p.T;                // rotate_vec_left(datas, 0, 1)  (mpc-util.hpp:17-22)
  const int slot = phys_t(p, p.T - 1);
  for (int b = 0; b < p.batch; b++) {
    const garb200_knot_view &k = shared ? knots[0] : knots[b];
    pack_knot(p.KL, k, s->h_cycle + (size_t)b * p.KL.stride);
    if (s->h_stage) {   // keep the staging mirror coherent: the appended knot is already on the device
      const size_t gs = (size_t)b * (p.T + 1) + slot;
      memcpy(s->h_stage + gs * p.KL.stride, s->h_cycle + (size_t)b * p.KL.stride, p.KL.stride * sizeof(double));
      s->dirty[gs] = 0;
    }
  }
  const size_t pitch_k = (size_t)(p.T + 1) * p.KL.stride * sizeof(double);
  CK(cudaMemcpy2DAsync(p.knots + (size_t)slot * p.KL.stride, pitch_k, s->h_cycle,
                       p.KL.stride * sizeof(double), p.KL.stride * sizeof(double), p.batch,
                       cudaMemcpyHostToDevice, s->stream));
  CK(cudaEventRecord(s->ev_h2d, s->stream));
  s->h2d_pending = true;
  // fresh StageFactor at N-1, kkt0 / thGrad / thHess reset (proximal-riccati.hxx:83-87)
  const size_t pitch_g = (size_t)(p.T + 1) * p.GL.stride * sizeof(double);
  CK(cudaMemset2DAsync(p.gains + (size_t)slot * p.GL.stride, pitch_g, 0, p.GL.stride * sizeof(double), p.batch, s->stream));
  const size_t B = p.batch;
  CK(cudaMemsetAsync(p.kkt0_ff, 0, std::max<size_t>(B * (p.nx + p.nc0), 1) * sizeof(double), s->stream));
  CK(cudaMemsetAsync(p.thGrad, 0, std::max<size_t>(B * p.nth, 1) * sizeof(double), s->stream));
  CK(cudaMemsetAsync(p.thHess, 0, std::max<size_t>(B * p.nth * p.nth, 1) * sizeof(double), s->stream));
  if (s->inner) {   // same ring offset and fresh StageFactor / kkt0 in the inner solver; user gains follow it
    garb200_solver *in = s->inner;
    DevParams &q = in->p;
    q.shift = p.shift;
    s->user_gains_fresh = false; // re-derived from the inner records (new offset, zeroed slot) on next use
    const size_t pq = (size_t)(q.T + 1) * q.GL.stride * sizeof(double);
    CK(cudaMemset2DAsync(q.gains + (size_t)slot * q.GL.stride, pq, 0, q.GL.stride * sizeof(double), q.batch, s->stream));
    CK(cudaMemsetAsync(q.kkt0_ff, 0, std::max<size_t>(B * (q.nx + q.nc0), 1) * sizeof(double), s->stream));
    CK(cudaMemsetAsync(q.thGrad, 0, std::max<size_t>(B * q.nth, 1) * sizeof(double), s->stream));
    CK(cudaMemsetAsync(q.thHess, 0, std::max<size_t>(B * q.nth * q.nth, 1) * sizeof(double), s->stream));
    in->did_forward = false; in->gains_fresh = false;
  }
  s->did_forward = false;
  s->gains_fresh = false;
  s->problem_dirty = s->inner_dirty = true;
  return GARB200_OK;
}

// ---- metric -------------------------------------------------------------------------------
int garb200_kkt_error(garb200_solver *s, double mudyn, double mueq, const double *theta, double *out) {
  NvtxRange nvtx_("garb200::kkt_error");
  if (!s || !out) return fail(GARB200_ERR_INVALID, "null argument");
  if (!s->did_forward) return fail(GARB200_ERR_STATE, "kkt_error needs a forward() into the solver-owned trajectory");
  CK(cudaSetDevice(s->cfg.device));
  const DevParams &p = s->p;
  const size_t B = p.batch;
  const double *d_theta = nullptr;
  if (theta && p.nth > 0 && p.num_legs < 2) {
    CK(cudaMemcpyAsync(p.theta, theta, B * p.nth * sizeof(double), cudaMemcpyHostToDevice, s->stream));
    d_theta = p.theta;
  }
  const int rc = launch_kkt_error(p, d_theta, mudyn, mueq, s->d_kkt_out, s->stream);
  if (rc < 0) return fail(rc == -1 ? GARB200_ERR_INVALID : GARB200_ERR_CUDA, "kkt_error launch failed");
  s->launches += rc;
  return fetch(s, out, s->d_kkt_out, 3 * B);
}

int garb200_upload_trajectory(garb200_solver *s, const double *xs, const double *us,
                              const double *vs, const double *lbdas) {
  if (!s || !xs || !lbdas) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(s->cfg.device));
  const DevParams &p = s->p;
  const size_t B = p.batch, n = p.nx, T = p.T;
  CK(cudaMemcpyAsync(p.xs, xs, B * (T + 1) * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  if (us && p.nu) CK(cudaMemcpyAsync(p.us, us, B * (T + 1) * p.nu * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  if (vs && p.nc) CK(cudaMemcpyAsync(p.vs, vs, B * (T + 1) * p.nc * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  CK(cudaMemcpyAsync(p.lbdas, lbdas, B * (p.nc0 + T * n) * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  s->did_forward = true;
  return GARB200_OK;
}

// ---- introspection --------------------------------------------------------------------------
int64_t garb200_kernel_launches(const garb200_solver *s) { return s ? s->launches : 0; }

int garb200_last_timings(garb200_solver *s, float *backward_ms, float *forward_ms) {
  if (!s) return fail(GARB200_ERR_INVALID, "null solver");
  CK(cudaSetDevice(s->cfg.device));
  CK(cudaStreamSynchronize(s->stream));
  if (backward_ms) { *backward_ms = 0.f; if (s->did_backward) CK(cudaEventElapsedTime(backward_ms, s->ev[0], s->ev[1])); }
  if (forward_ms) { *forward_ms = 0.f; if (s->did_forward) CK(cudaEventElapsedTime(forward_ms, s->ev[2], s->ev[3])); }
  return GARB200_OK;
}

int garb200_used_fast_path(const garb200_solver *s) { return s && s->used_fast ? 1 : 0; }
int garb200_used_explicit_path(const garb200_solver *s) { return s && s->used_fast && s->p.explicit_dyn ? 1 : 0; }

int garb200_phase_cycles(garb200_solver *s, long long *out16) {
  if (!s || !out16) return fail(GARB200_ERR_INVALID, "null argument");
  if (s->inner) return garb200_phase_cycles(s->inner, out16);
  if (!s->p.phase_cycles) return fail(GARB200_ERR_STATE, "create the solver with GARB200_PHASE_TIMERS in garb200_config.flags");
  CK(cudaSetDevice(s->cfg.device));
  CK(cudaStreamSynchronize(s->stream));
  CK(cudaMemcpy(out16, s->p.phase_cycles, 16 * sizeof(long long), cudaMemcpyDeviceToHost));
  CK(cudaMemset(s->p.phase_cycles, 0, 16 * sizeof(long long)));
  return GARB200_OK;
}

int garb200_timing_log_start(garb200_solver *s, int capacity) {
  if (!s || capacity < 0) return fail(GARB200_ERR_INVALID, "bad argument");
  CK(cudaSetDevice(s->cfg.device));
  while ((int)s->tlog.size() < 4 * capacity) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    s->tlog.push_back(e);
  }
  s->tlog_cap = capacity; s->tlog_bwd = s->tlog_fwd = 0;
  return GARB200_OK;
}

int garb200_timing_log_read(garb200_solver *s, float *backward_ms, float *forward_ms, int *count) {
  if (!s || !count) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(s->cfg.device));
  CK(cudaStreamSynchronize(s->stream));
  const int nb = s->tlog_bwd, nf = s->tlog_fwd;
  for (int i = 0; i < nb && backward_ms; i++) CK(cudaEventElapsedTime(backward_ms + i, s->tlog[4 * i], s->tlog[4 * i + 1]));
  for (int i = 0; i < nf && forward_ms; i++) CK(cudaEventElapsedTime(forward_ms + i, s->tlog[4 * i + 2], s->tlog[4 * i + 3]));
  *count = nb < nf ? nb : nf;
  s->tlog_cap = 0;
  return GARB200_OK;
}

int garb200_fp64_peak(int device, int kind, double *tflops) {
  if (!tflops) return fail(GARB200_ERR_INVALID, "null argument");
  CK(cudaSetDevice(device));
  const int rc = fp64_peak(kind, tflops);
  return rc == 0 ? GARB200_OK : fail(GARB200_ERR_CUDA, "fp64 micro-benchmark failed");
}

} // extern "C"
