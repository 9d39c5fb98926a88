// forward.cu — batched forward rollout and KKT-residual kernels (sm_100a).
//
//   forward_kernel    ProximalRiccatiKernel::computeInitial + forwardImpl
//                     riccati-kernel.hxx:153-164, 303-361; ProximalRiccatiSolver::forward
//                     proximal-riccati.hxx:66-78; ParallelRiccatiSolver::forward
//                     parallel-solver.hxx:181-215 (one CTA per leg)
//   kkt_error_kernel  lqrComputeKktError  utils.hxx:94-192
//
// The rollout is a chain of GEMVs over the gains records written by the backward sweep:
// HBM-bound (reads (nu+nc+2nx)(nx+1+nth) doubles per knot, 2 flops each).  Gains records
// are streamed with 1-D TMA bulk copies through an NSTAGE-deep shared-memory ring so the
// loads of knots t+1.. are in flight while knot t is applied; 4 threads share one row of
// the row-major gain matrix (bank-conflict free for nx = 4 mod 16, e.g. 12, 36, 76).
#include "device_linalg.cuh"
#include <cstdlib>
#include "kernels.cuh"

namespace garb200 {

template <int NT, int NSTAGE>
__global__ void __launch_bounds__(NT) forward_kernel(const DevParams p, const double *theta_in,
                                                     double *xs, double *us, double *vs,
                                                     double *lbdas) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.nx, T = p.T, c0 = p.nc0;
  const int nlegs = p.num_legs >= 2 ? p.num_legs : 1;
  const bool legmode = nlegs > 1;
  const int b = blockIdx.x / nlegs, leg = blockIdx.x % nlegs;
  int beg = 0, end = T + 1;
  if (legmode) get_work(T, leg, nlegs, beg, end);
  const int h = legmode ? (leg < nlegs - 1 ? p.hleg : 0) : (theta_in ? p.nth : 0);
  const int hl = p.GL.nth;  // row stride of fth in the record
  const int gstride = p.GL.stride;
  const int tid = threadIdx.x;

  double *ring = sm;                                   // NSTAGE * gstride
  double *x = ring + (size_t)NSTAGE * gstride;         // n
  double *th = x + ((n + 1) & ~1);                     // max(h,1)
  double *outv = th + ((p.GL.nth + 2) & ~1);           // nrows
  uint64_t *bars = reinterpret_cast<uint64_t *>(outv + ((p.GL.nrows + 1) & ~1));

  const size_t xoff = (size_t)b * (T + 1) * n, uoff = (size_t)b * (T + 1) * p.nu,
               voff = (size_t)b * (T + 1) * p.nc, loff = (size_t)b * (c0 + (size_t)T * n);
  auto lam = [&](int t) { return lbdas + loff + (t == 0 ? 0 : c0 + (size_t)(t - 1) * n); };

  if (tid == 0) {
    for (int s = 0; s < NSTAGE; s++) mbar_init(bars + s, 1);
    fence_mbar_init();
  }
  // initial state / parameter of this sweep
  if (!legmode) {
    const int s0 = n + c0;
    const double *ff0 = p.kkt0_ff + (size_t)b * s0;
    const double *fth0 = p.kkt0_fth + (size_t)b * s0 * p.nth;
    for (int i = tid; i < h; i += NT) th[i] = theta_in[(size_t)b * p.nth + i];
    __syncthreads();
    for (int i = tid; i < s0; i += NT) {            // computeInitial, riccati-kernel.hxx:153-164
      double v = ff0[i];
      for (int j = 0; j < h; j++) v = fma(fth0[i * p.nth + j], th[j], v);
      if (i < n) { x[i] = v; xs[xoff + i] = v; }
      else lam(0)[i - n] = v;
    }
  } else {
    // condensed solution: lbdas[i0] = sol[2 leg], xs[i0] = sol[2 leg + 1]   (parallel-solver.hxx:188-192)
    const double *sol = p.csol + (size_t)b * (c0 + (size_t)(2 * nlegs - 1) * n);
    auto blk = [&](int k) { return sol + (k == 0 ? 0 : c0 + (size_t)(k - 1) * n); };
    const int nl = leg == 0 ? c0 : n;
    for (int i = tid; i < nl; i += NT) lam(beg)[i] = blk(2 * leg)[i];
    for (int i = tid; i < n; i += NT) { const double v = blk(2 * leg + 1)[i]; x[i] = v; xs[xoff + (size_t)beg * n + i] = v; }
    for (int i = tid; i < h; i += NT) th[i] = blk(2 * (leg + 1))[i];   // theta = lbdas[end]
  }
  __syncthreads();

  const size_t gbase = (size_t)b * (T + 1);
  const uint32_t gbytes = (uint32_t)gstride * 8u;
  const int nk = end - beg;
  if (tid == 0) {
    for (int s = 0; s < NSTAGE && s < nk; s++) {
      mbar_expect_tx(bars + s, gbytes);
      tma_load_1d(ring + (size_t)s * gstride, p.gains + (gbase + phys_t(p, beg + s)) * gstride, gbytes, bars + s);
    }
  }
  const int q = tid & 3, rl = tid >> 2;   // 4 threads per gain row
  for (int k = 0; k < nk; k++) {
    const int t = beg + k, s = k % NSTAGE;
    mbar_wait(bars + s, (uint32_t)((k / NSTAGE) & 1));
    const double *rec = ring + (size_t)s * gstride;
    const double *ff = rec + p.GL.ff, *fb = rec + p.GL.fb, *fth = rec + p.GL.fth;
    const int m = p.nu_t[t], c = p.nc_t[t];
    const bool last = (t == end - 1);
    const int nr = last ? m + c : m + c + 2 * n;
    for (int r0 = 0; r0 < nr; r0 += NT / 4) {
      const int r = r0 + rl;
      double acc = 0.0;
      if (r < nr) {
        const double *row = fb + (size_t)r * n;
        for (int j = q; j < n; j += 4) acc = fma(row[j], x[j], acc);
        const double *rowt = fth + (size_t)r * hl;
        for (int j = q; j < h; j += 4) acc = fma(rowt[j], th[j], acc);
      }
      acc += __shfl_xor_sync(0xffffffffu, acc, 1);
      acc += __shfl_xor_sync(0xffffffffu, acc, 2);
      if (r < nr && q == 0) outv[r] = ff[r] + acc;
    }
    __syncthreads();   // outv complete, all reads of x and of ring slot s done
    if (tid == 0 && k + NSTAGE < nk) {
      mbar_expect_tx(bars + s, gbytes);
      tma_load_1d(ring + (size_t)s * gstride, p.gains + (gbase + phys_t(p, t + NSTAGE)) * gstride, gbytes, bars + s);
    }
    for (int i = tid; i < nr; i += NT) {
      const double v = outv[i];
      if (i < m) us[uoff + (size_t)t * p.nu + i] = v;
      else if (i < m + c) vs[voff + (size_t)t * p.nc + (i - m)] = v;
      else if (i < m + c + n) lam(t + 1)[i - m - c] = v;
      else { x[i - m - c - n] = v; xs[xoff + (size_t)(t + 1) * n + (i - m - c - n)] = v; }
    }
    // zero padding of ragged knots so the packed arrays are fully defined
    for (int i = m + tid; i < p.nu; i += NT) us[uoff + (size_t)t * p.nu + i] = 0.0;
    for (int i = c + tid; i < p.nc; i += NT) vs[voff + (size_t)t * p.nc + i] = 0.0;
    __syncthreads();
  }
}

// ---- small-record rollout: one WARP per problem, one lane per gain row -------------------
// For records with nu + 2 nx <= 32 rows (the nx = 12 shape), nc = nth = 0, uniform knots: every
// lane owns one row of [K; L; A] and keeps x in registers, so a knot is 6 vector loads + NX FMAs
// per lane, one shared-memory broadcast of the new state and ONE store per lane — about 40
// warp instructions per knot instead of ~500 for the runtime-dimension kernel above.  WPB
// warps per CTA, each with its own NSTAGE-deep TMA ring and mbarriers; no CTA barrier at all.
template <int NX, int NU, int WPB, int NSTAGE>
__global__ void __launch_bounds__(32 * WPB) forward_small_kernel(const DevParams p, double *xs, double *us,
                                                                   double *lbdas) {
  constexpr int n = NX, m = NU, nrows = NU + 2 * NX;
  static_assert(nrows <= 32 && NX % 2 == 0, "one lane per gain row");
  constexpr int gff = 0, gfb = (nrows + 1) & ~1, gstride = (gfb + nrows * n + 1) & ~1;
  extern __shared__ __align__(16) double sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * WPB + warp;
  if (b >= p.batch) return;
  const int T = p.T, c0 = p.nc0;
  double *ring = sm + (size_t)warp * (NSTAGE * gstride + n + NSTAGE);
  double *xsh = ring + NSTAGE * gstride;
  uint64_t *bars = reinterpret_cast<uint64_t *>(xsh + n);
  const size_t gbase = (size_t)b * (T + 1);
  if (lane == 0) {
    for (int s = 0; s < NSTAGE; s++) mbar_init(bars + s, 1);
    fence_mbar_init();
  }
  __syncwarp();
  if (lane == 0)
    for (int s = 0; s < NSTAGE && s < T; s++) {
      mbar_expect_tx(bars + s, gstride * 8u);
      tma_load_1d(ring + s * gstride, p.gains + (gbase + phys_t(p, s)) * gstride, gstride * 8u, bars + s);
    }
  // computeInitial (riccati-kernel.hxx:153-164): x0 = kkt0.ff[0:n], lbd0 = kkt0.ff[n:n+c0]
  const double *ff0 = p.kkt0_ff + (size_t)b * (n + c0);
  const size_t xoff = (size_t)b * (T + 1) * n, uoff = (size_t)b * (T + 1) * m, loff = (size_t)b * (c0 + (size_t)T * n);
  double x[n];
#pragma unroll
  for (int j = 0; j < n; j++) x[j] = ff0[j];
  if (lane < n) xs[xoff + lane] = ff0[lane];
  for (int i = lane; i < c0; i += 32) lbdas[loff + i] = ff0[n + i];
  // where this lane's row goes: u (rows < m), lbd_{t+1} (next n rows), x_{t+1} (last n rows)
  double *dst = nullptr; size_t dstride = 0;
  if (lane < m) { dst = us + uoff + lane; dstride = m; }
  else if (lane < m + n) { dst = lbdas + loff + c0 + (lane - m); dstride = n; }
  else { dst = xs + xoff + n + (lane - m - n); dstride = n; }
  const bool active = lane < nrows;
  const int row = active ? lane : nrows - 1;
#pragma unroll 1
  for (int t = 0; t < T; t++) {
    const int s = t % NSTAGE;
    mbar_wait(bars + s, (uint32_t)((t / NSTAGE) & 1));
    const double *rec = ring + s * gstride;
    const double2 *rw = reinterpret_cast<const double2 *>(rec + gfb + row * n);
    double acc = rec[gff + row];
#pragma unroll
    for (int j = 0; j < n / 2; j++) { const double2 v = rw[j]; acc = fma(v.x, x[2 * j], acc); acc = fma(v.y, x[2 * j + 1], acc); }
    if (active) dst[(size_t)t * dstride] = acc;
    if (lane >= m + n && active) xsh[lane - m - n] = acc;
    __syncwarp();                       // xsh complete; every lane is done with ring slot s
    if (lane == 0 && t + NSTAGE < T) {
      fence_proxy_async();
      mbar_expect_tx(bars + s, gstride * 8u);
      tma_load_1d(ring + s * gstride, p.gains + (gbase + phys_t(p, t + NSTAGE)) * gstride, gstride * 8u, bars + s);
    }
#pragma unroll
    for (int j = 0; j < n / 2; j++) { const double2 v = reinterpret_cast<const double2 *>(xsh)[j]; x[2 * j] = v.x; x[2 * j + 1] = v.y; }
    __syncwarp();
  }
  if (lane < m) us[uoff + (size_t)T * m + lane] = 0.0;     // terminal knot has no control: defined padding
}

template <int NX, int NU>
static int launch_forward_small(const DevParams &p, double *xs, double *us, double *lbdas, cudaStream_t st) {
  constexpr int WPB = 4, NSTAGE = 4;
  constexpr int nrows = NU + 2 * NX, gfb = (nrows + 1) & ~1, gstride = (gfb + nrows * NX + 1) & ~1;
  if (p.GL.stride != gstride || p.GL.fb != gfb) return -3;
  const size_t smem = (size_t)WPB * (NSTAGE * gstride + NX + NSTAGE) * sizeof(double);
  auto kern = forward_small_kernel<NX, NU, WPB, NSTAGE>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -2;
  kern<<<(p.batch + WPB - 1) / WPB, 32 * WPB, smem, st>>>(p, xs, us, lbdas);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

static size_t forward_smem(const DevParams &p, int nstage) {
  size_t d = (size_t)nstage * p.GL.stride + ((p.nx + 1) & ~1) + ((p.GL.nth + 2) & ~1) +
             ((p.GL.nrows + 1) & ~1) + nstage + 2;
  return d * sizeof(double);
}

template <int NT, int NSTAGE>
static int launch_forward_t(const DevParams &p, const double *d_theta, double *xs, double *us,
                            double *vs, double *lbdas, cudaStream_t st) {
  const size_t smem = forward_smem(p, NSTAGE);
  if (smem > 227u * 1024u) return -1;
  auto kern = forward_kernel<NT, NSTAGE>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return -2;
  const int nlegs = p.num_legs >= 2 ? p.num_legs : 1;
  kern<<<p.batch * nlegs, NT, smem, st>>>(p, d_theta, xs, us, vs, lbdas);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

int launch_forward(const DevParams &p, const double *d_theta, double *xs, double *us, double *vs,
                   double *lbdas, cudaStream_t st) {
  if (p.uniform_knots && p.num_legs < 2 && p.nc == 0 && p.GL.nth == 0 && p.nx == 12 && p.nu == 4) {
    const int rc = launch_forward_small<12, 4>(p, xs, us, lbdas, st);
    if (rc != -3) return rc;
  }
  if (p.GL.stride * 8 <= 4096) return launch_forward_t<32, 4>(p, d_theta, xs, us, vs, lbdas, st);
  const int stages = p.tune.fwd_stages;   // 2: more CTAs per SM beats a deeper ring (c1b 5.7 -> 6.6 TB/s, c3 4.1 -> 5.5 TB/s)
  if (stages == 1 || (stages == 2 && forward_smem(p, 2) > 114u * 1024u && forward_smem(p, 1) <= 113u * 1024u + 512u))   // two 1-deep CTAs per SM
    return launch_forward_t<128, 1>(p, d_theta, xs, us, vs, lbdas, st);
  if (stages == 4 && forward_smem(p, 4) <= 110u * 1024u) return launch_forward_t<128, 4>(p, d_theta, xs, us, vs, lbdas, st);
  if (stages >= 3 && forward_smem(p, 3) <= 100u * 1024u) return launch_forward_t<128, 3>(p, d_theta, xs, us, vs, lbdas, st);
  return launch_forward_t<128, 2>(p, d_theta, xs, us, vs, lbdas, st);
}

// ---- lqrComputeKktError (utils.hxx:94-192): one CTA per problem, metric only ------------
__device__ __forceinline__ double nanmax(double cur, double v) {
  const double a = fabs(v);
  return (!(a <= cur)) ? a : cur;   // NaN propagates like Eigen's lpNorm<Infinity>
}

template <int NT>
__global__ void __launch_bounds__(NT) kkt_error_kernel(const DevParams p, const double *theta_in,
                                                       double mudyn, double mueq,
                                                       const double *xs, const double *us,
                                                       const double *vs, const double *lbdas,
                                                       double *out) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.nx, T = p.T, c0 = p.nc0, tid = threadIdx.x, b = blockIdx.x;
  const garb200_knot_layout &KL = p.KL;
  const int h = theta_in ? p.nth : 0;
  const double *Eprev = nullptr;   // E of knot t-1, read straight from global (metric kernel)
  double *x = sm, *xn = x + n, *u = xn + n, *v = u + p.nu + 1, *lam = v + p.nc + 1;
  double *lamn = lam + (n > c0 ? n : c0), *th = lamn + n, *red = th + p.nth + 1;
  const size_t xoff = (size_t)b * (T + 1) * n, uoff = (size_t)b * (T + 1) * p.nu,
               voff = (size_t)b * (T + 1) * p.nc, loff = (size_t)b * (c0 + (size_t)T * n);
  auto lamp = [&](int t) { return lbdas + loff + (t == 0 ? 0 : c0 + (size_t)(t - 1) * n); };
  double dynE = 0.0, cstE = 0.0, dualE = 0.0;
  for (int i = tid; i < h; i += NT) th[i] = theta_in[(size_t)b * p.nth + i];
  {
    const double *G0 = p.G0 + (size_t)b * c0 * n, *g0 = p.g0 + (size_t)b * c0;
    for (int i = tid; i < c0; i += NT) {                 // :122-126
      double r = g0[i];
      for (int j = 0; j < n; j++) r = fma(G0[i + j * c0], xs[xoff + j], r);
      r -= mudyn * lamp(0)[i];
      dynE = nanmax(dynE, r);
    }
  }
  for (int t = 0; t <= T; t++) {
    const int m = p.nu_t[t], c = p.nc_t[t];
    const double *grec = p.knots + ((size_t)b * (T + 1) + phys_t(p, t)) * KL.stride;
    const double *rec = grec;
    __syncthreads();
    for (int i = tid; i < n; i += NT) {
      x[i] = xs[xoff + (size_t)t * n + i];
      if (t < T) { xn[i] = xs[xoff + (size_t)(t + 1) * n + i]; lamn[i] = lamp(t + 1)[i]; }
    }
    for (int i = tid; i < (t == 0 ? c0 : n); i += NT) lam[i] = lamp(t)[i];
    for (int i = tid; i < m; i += NT) u[i] = us[uoff + (size_t)t * p.nu + i];
    for (int i = tid; i < c; i += NT) v[i] = vs[voff + (size_t)t * p.nc + i];
    __syncthreads();
    const double *Q = rec + KL.Q, *S = rec + KL.S, *R = rec + KL.R, *q = rec + KL.q, *r = rec + KL.r;
    const double *A = rec + KL.A, *B = rec + KL.B, *E = rec + KL.E, *f = rec + KL.f;
    const double *C = rec + KL.C, *D = rec + KL.D, *d = rec + KL.d;
    const double *Gx = rec + KL.Gx, *Gu = rec + KL.Gu;
    const double *G0 = p.G0 + (size_t)b * c0 * n;
    for (int i = tid; i < c; i += NT) {                    // :143, :148
      double a = d[i];
      for (int j = 0; j < n; j++) a = fma(C[i + j * c], x[j], a);
      a -= mueq * v[i];
      for (int j = 0; j < m; j++) a = fma(D[i + j * c], u[j], a);
      cstE = nanmax(cstE, a);
    }
    for (int i = tid; i < n; i += NT) {                    // gx :144, :149, :153-158, :165, :175
      double a = q[i];
      for (int j = 0; j < n; j++) a = fma(Q[i + j * n], x[j], a);
      for (int k = 0; k < c; k++) a = fma(C[k + i * c], v[k], a);
      for (int j = 0; j < m; j++) a = fma(S[i + j * n], u[j], a);
      if (t == 0) { for (int k = 0; k < c0; k++) a = fma(G0[k + i * c0], lam[k], a); }
      else { for (int k = 0; k < n; k++) a = fma(Eprev[k + i * n], lam[k], a); }
      if (t < T) for (int k = 0; k < n; k++) a = fma(A[k + i * n], lamn[k], a);
      for (int j = 0; j < h; j++) a = fma(Gx[i + j * n], th[j], a);
      dualE = nanmax(dualE, a);
    }
    for (int i = tid; i < m; i += NT) {                    // gu :145, :150, :166, :176
      double a = r[i];
      for (int j = 0; j < n; j++) a = fma(S[j + i * n], x[j], a);
      for (int k = 0; k < c; k++) a = fma(D[k + i * c], v[k], a);
      for (int j = 0; j < m; j++) a = fma(R[i + j * m], u[j], a);
      if (t < T) for (int k = 0; k < n; k++) a = fma(B[k + i * n], lamn[k], a);
      for (int j = 0; j < h; j++) a = fma(Gu[i + j * m], th[j], a);
      dualE = nanmax(dualE, a);
    }
    if (t < T) {
      for (int i = tid; i < n; i += NT) {                  // :161-163
        double a = f[i];
        for (int j = 0; j < n; j++) a = fma(A[i + j * n], x[j], a);
        for (int j = 0; j < m; j++) a = fma(B[i + j * n], u[j], a);
        for (int j = 0; j < n; j++) a = fma(E[i + j * n], xn[j], a);
        a -= mudyn * lamn[i];
        dynE = nanmax(dynE, a);
      }
    }
    Eprev = E;
  }
  red[tid * 3 + 0] = dynE; red[tid * 3 + 1] = cstE; red[tid * 3 + 2] = dualE;
  __syncthreads();
  if (tid == 0) {
    double a = 0.0, c = 0.0, d = 0.0;
    for (int i = 0; i < NT; i++) { a = nanmax(a, red[i * 3]); c = nanmax(c, red[i * 3 + 1]); d = nanmax(d, red[i * 3 + 2]); }
    out[(size_t)b * 3 + 0] = a; out[(size_t)b * 3 + 1] = c; out[(size_t)b * 3 + 2] = d;
  }
}

int launch_kkt_error(const DevParams &p, const double *d_theta, double mudyn, double mueq,
                     double *d_out, cudaStream_t st) {
  constexpr int NT = 128;
  const int n = p.nx;
  const size_t d = (size_t)4 * n + (n > p.nc0 ? n : p.nc0) + p.nu + p.nc + p.nth + 8 + 3 * NT;
  const size_t smem = d * sizeof(double);
  if (smem > 227u * 1024u) return -1;
  auto kern = kkt_error_kernel<NT>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return -2;
  kern<<<p.batch, NT, smem, st>>>(p, d_theta, mudyn, mueq, p.xs, p.us, p.vs, p.lbdas, d_out);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

} // namespace garb200
