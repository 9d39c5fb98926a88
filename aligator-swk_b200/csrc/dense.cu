// dense.cu — RiccatiSolverDense on the GPU: the reference's dense per-knot KKT solver
//   DenseKernel::terminalSolve / stageKernelSolve      gar/dense-kernel.hpp:56-95, 97-172
//   RiccatiSolverDense::backward (incl. initial stage) gar/dense-riccati.hxx:46-94
// One Bunch-Kaufman factorization (bunchkaufman.hpp:20-167, CTA-parallel restatement bk_factor_cta) of the
// (nu + nc + 2 nx)^2 stage KKT matrix per knot
//   [ R   D^T   B^T    0   ]   unknown block rows [K; Z; L; Y]: the same rows as the proximal solver's
//   [ D  -mueq  0      0   ]   [K; Z; L; A] gains record, so the rollout kernel (forward.cu) and every gains
//   [ B   0    -mudyn  E   ]   reader are shared.  It needs no E^{-1}, no Schur complement and no inertia
//   [ 0   0     E^T    Pxx+]   assumption: the robustness fallback for badly conditioned / indefinite stages.
// Unlike the proximal kernel it READS Gv (dense-kernel.hpp:76,134); Gv is not part of the packed knot record
// and comes in a separate optional array (garb200_upload_gv).
// Runtime dimensions, one CTA per problem, knot data read straight from global memory: this is the slow,
// sturdy path (selected with GARB200_DENSE_KERNEL), not a speed path.
#include "device_linalg.cuh"
#include "kernels.cuh"

namespace garb200 {
namespace {

__host__ __device__ inline int dr2(int x) { return (x + 1) & ~1; }

struct DensePlan {
  int nrmax, ld, ncol, vsz;
  int K, R, Va, Vb, wk, sub, ints, total;   // offsets in doubles
};
__host__ __device__ inline DensePlan dense_plan(const DevParams &p) {
  DensePlan d;
  const int n = p.nx, h = p.nth;
  const int a = p.nu + p.nc + 2 * n, b = n + p.nc0;
  d.nrmax = a > b ? a : b;
  d.ld = d.nrmax | 1;
  d.ncol = 1 + n + h;
  d.vsz = dr2(n * n) + dr2(n) + dr2(n * h) + dr2(h * h) + dr2(h);
  int o = 0;
  d.K = o;    o = dr2(o + d.nrmax * d.ld);
  d.R = o;    o = dr2(o + d.nrmax * d.ncol);
  d.Va = o;   o = dr2(o + d.vsz);
  d.Vb = o;   o = dr2(o + d.vsz);
  d.wk = o;   o = dr2(o + 2 * d.nrmax + 4);
  d.sub = o;  o = dr2(o + d.nrmax + 2);
  d.ints = o; o = dr2(o + (d.nrmax + 8 + 1) / 2);
  d.total = o;
  return d;
}

struct ValuePtr { double *Pxx, *px, *Pxt, *Ptt, *pt; };
__device__ inline ValuePtr value_ptr(double *base, int n, int h) {
  ValuePtr v;
  v.Pxx = base; v.px = v.Pxx + dr2(n * n); v.Pxt = v.px + dr2(n); v.Ptt = v.Pxt + dr2(n * h); v.pt = v.Ptt + dr2(h * h);
  return v;
}

template <int NT>
__global__ void __launch_bounds__(NT) dense_backward_kernel(const DevParams p) {
  extern __shared__ __align__(16) double sm[];
  const DensePlan pl = dense_plan(p);
  const int n = p.nx, h = p.nth, T = p.T, c0 = p.nc0, tid = threadIdx.x, b = blockIdx.x;
  const int ld = pl.ld, nrm = pl.nrmax;
  const double mudyn = p.mudyn, mueq = p.mueq;
  double *Km = sm + pl.K, *Rh = sm + pl.R, *wk = sm + pl.wk, *sub = sm + pl.sub;
  int *piv = reinterpret_cast<int *>(sm + pl.ints), *ctl = piv + nrm + 2;
  ValuePtr cur = value_ptr(sm + pl.Va, n, h), nxt = value_ptr(sm + pl.Vb, n, h);
  const garb200_knot_layout &KL = p.KL;
  int flags = 0;
  auto rhs = [&](int i, int j) -> double & { return Rh[i + j * nrm]; };

  for (int t = T; t >= 0; t--) {
    const int m = p.nu_t[t], c = p.nc_t[t];
    const bool term = (t == T);
    const int nr = term ? m + c : m + c + 2 * n;
    const int oZ = m, oL = m + c, oY = m + c + n;
    const double *rec = p.knots + ((size_t)b * (T + 1) + phys_t(p, t)) * KL.stride;
    const double *Q = rec + KL.Q, *S = rec + KL.S, *R = rec + KL.R, *q = rec + KL.q, *r = rec + KL.r;
    const double *A = rec + KL.A, *B = rec + KL.B, *E = rec + KL.E, *f = rec + KL.f;
    const double *C = rec + KL.C, *D = rec + KL.D, *d = rec + KL.d;
    const double *Gth = rec + KL.Gth, *Gx = rec + KL.Gx, *Gu = rec + KL.Gu, *gam = rec + KL.gamma;
    const double *Gv = p.gv ? p.gv + ((size_t)b * (T + 1) + t) * (size_t)(p.nc * h) : nullptr;   // ld = p.nc
    auto gvv = [&](int i, int a) { return Gv ? __ldg(Gv + i + a * p.nc) : 0.0; };
    // ---- KKT matrix (lower triangle is what Bunch-Kaufman reads) and right-hand sides   (:56-64, 99-112)
    for (int e = tid; e < nr * ld; e += NT) Km[e] = 0.0;
    for (int e = tid; e < nr * pl.ncol; e += NT) Rh[(e % nr) + (e / nr) * nrm] = 0.0;
    __syncthreads();
    for (int e = tid; e < m * m; e += NT) Km[(e % m) + (e / m) * ld] = __ldg(R + e);
    for (int e = tid; e < c * m; e += NT) Km[(oZ + e % c) + (e / c) * ld] = __ldg(D + e);
    for (int i = tid; i < c; i += NT) Km[(oZ + i) + (oZ + i) * ld] = -mueq;
    for (int i = tid; i < m; i += NT) rhs(i, 0) = -__ldg(r + i);
    for (int i = tid; i < c; i += NT) rhs(oZ + i, 0) = -__ldg(d + i);
    for (int e = tid; e < n * m; e += NT) rhs(e / n, 1 + e % n) = -__ldg(S + e);            // K rows: -S^T
    for (int e = tid; e < c * n; e += NT) rhs(oZ + e % c, 1 + e / c) = -__ldg(C + e);       // Z rows: -C
    for (int e = tid; e < m * h; e += NT) rhs(e % m, 1 + n + e / m) = -__ldg(Gu + e);       // Kth rows: -Gu
    for (int e = tid; e < c * h; e += NT) rhs(oZ + e % c, 1 + n + e / c) = -gvv(e % c, e / c);   // Zth rows: -Gv
    if (!term) {
      for (int e = tid; e < n * m; e += NT) Km[(oL + e % n) + (e / n) * ld] = __ldg(B + e);
      for (int i = tid; i < n; i += NT) Km[(oL + i) + (oL + i) * ld] = -mudyn;
      for (int e = tid; e < n * n; e += NT) Km[(oY + e / n) + (oL + e % n) * ld] = __ldg(E + e);   // E^T below the diagonal
      for (int e = tid; e < n * n; e += NT) Km[(oY + e % n) + (oY + e / n) * ld] = nxt.Pxx[e];
      for (int i = tid; i < n; i += NT) { rhs(oL + i, 0) = -__ldg(f + i); rhs(oY + i, 0) = -nxt.px[i]; }
      for (int e = tid; e < n * n; e += NT) rhs(oL + e % n, 1 + e / n) = -__ldg(A + e);      // L rows: -A
      for (int e = tid; e < n * h; e += NT) rhs(oY + e % n, 1 + n + e / n) = -nxt.Pxt[e];    // Yth rows: -Pxt+
    }
    __syncthreads();
    // ---- factorize and solve                                                              (:66-75, 115-141)
    if (nr > 0) {
      if (!bk_factor_cta<NT>(tid, nr, Km, ld, piv, sub, wk, ctl)) flags |= GARB200_FLAG_BK_SINGULAR;
      __syncthreads();
      for (int col = tid; col < pl.ncol; col += NT) bk_solve_column(nr, Km, ld, piv, sub, Rh + (size_t)col * nrm, 1);
      __syncthreads();
    }
    // ---- value function of this knot                                                      (:77-94, 143-171)
    auto Kf = [&](int k, int j) { return rhs(k, 1 + j); };          // fb(k, j), k over all block rows
    auto Tf = [&](int k, int a) { return rhs(k, 1 + n + a); };      // ft(k, a)
    for (int e = tid; e < n * n; e += NT) {
      const int i = e % n, j = e / n;
      double acc = __ldg(Q + e);
      for (int k = 0; k < m; k++) acc = fma(__ldg(S + i + k * n), Kf(k, j), acc);
      for (int k = 0; k < c; k++) acc = fma(__ldg(C + k + i * c), Kf(oZ + k, j), acc);
      if (!term) for (int k = 0; k < n; k++) acc = fma(__ldg(A + k + i * n), Kf(oL + k, j), acc);
      cur.Pxx[e] = acc;
    }
    for (int i = tid; i < n; i += NT) {
      double acc = __ldg(q + i);
      for (int k = 0; k < m; k++) acc = fma(__ldg(S + i + k * n), rhs(k, 0), acc);
      for (int k = 0; k < c; k++) acc = fma(__ldg(C + k + i * c), rhs(oZ + k, 0), acc);
      if (!term) for (int k = 0; k < n; k++) acc = fma(__ldg(A + k + i * n), rhs(oL + k, 0), acc);
      cur.px[i] = acc;
    }
    for (int e = tid; e < n * h; e += NT) {
      const int i = e % n, a = e / n;
      double acc = __ldg(Gx + e);
      for (int k = 0; k < m; k++) acc = fma(Kf(k, i), __ldg(Gu + k + a * m), acc);
      for (int k = 0; k < c; k++) acc = fma(Kf(oZ + k, i), gvv(k, a), acc);
      if (!term) for (int k = 0; k < n; k++) acc = fma(Kf(oY + k, i), nxt.Pxt[k + a * n], acc);
      cur.Pxt[e] = acc;
    }
    for (int e = tid; e < h * h + h; e += NT) {
      if (e < h * h) {
        const int a = e % h, bb = e / h;
        double acc = __ldg(Gth + e);
        for (int k = 0; k < m; k++) acc = fma(Tf(k, a), __ldg(Gu + k + bb * m), acc);
        for (int k = 0; k < c; k++) acc = fma(Tf(oZ + k, a), gvv(k, bb), acc);
        if (!term) for (int k = 0; k < n; k++) acc = fma(Tf(oY + k, a), nxt.Pxt[k + bb * n], acc);
        cur.Ptt[e] = acc;
      } else {
        const int a = e - h * h;
        double acc = __ldg(gam + a);
        for (int k = 0; k < m; k++) acc = fma(__ldg(Gu + k + a * m), rhs(k, 0), acc);
        for (int k = 0; k < c; k++) acc = fma(gvv(k, a), rhs(oZ + k, 0), acc);
        if (!term) for (int k = 0; k < n; k++) acc = fma(nxt.Pxt[k + a * n], rhs(oY + k, 0), acc);
        cur.pt[a] = acc;
      }
    }
    // ---- gains record [ff | fb | fth] (row-major fb / fth, riccati-kernel.hpp:76-78 layout) -> HBM
    double *g = p.gains + ((size_t)b * (T + 1) + phys_t(p, t)) * p.GL.stride;
    const int hl = p.GL.nth;
    for (int e = tid; e < p.GL.stride; e += NT) {
      double v = 0.0;
      if (e >= p.GL.fth) { const int k = e - p.GL.fth, i = hl ? k / hl : 0, a = hl ? k % hl : 0; if (hl && i < nr && a < h) v = Tf(i, a); }
      else if (e >= p.GL.fb) { const int k = e - p.GL.fb, i = k / n, j = k % n; if (i < nr) v = Kf(i, j); }
      else if (e < nr) v = rhs(e, 0);
      g[e] = v;
    }
    __syncthreads();                                    // value complete, all reads of the solved block done
    const int slot = p.store_all_values ? t : (t == 0 ? 0 : -1);
    if (slot >= 0) {
      double *v = p.values + ((size_t)b * p.nslots + slot) * p.VL.stride;
      for (int e = tid; e < n * n; e += NT) v[p.VL.Pmat + e] = cur.Pxx[e];
      for (int i = tid; i < n; i += NT) v[p.VL.pvec + i] = cur.px[i];
      for (int e = tid; e < n * h; e += NT) v[p.VL.Vxt + e] = cur.Pxt[e];
      for (int e = tid; e < h * h; e += NT) v[p.VL.Vtt + e] = cur.Ptt[e];
      for (int i = tid; i < h; i += NT) v[p.VL.vt + i] = cur.pt[i];
    }
    const ValuePtr tmp = cur; cur = nxt; nxt = tmp;     // this knot's value is the next knot's (t - 1) "vn"
  }
  // ---- initial stage: BK of [Pxx0 G0^T; G0 -mudyn I]                              (dense-riccati.hxx:68-92)
  {
    const ValuePtr v0 = nxt;
    const int s0 = n + c0;
    const double *G0 = p.G0 + (size_t)b * c0 * n, *g0 = p.g0 + (size_t)b * c0;
    for (int e = tid; e < s0 * ld; e += NT) Km[e] = 0.0;
    for (int e = tid; e < s0 * (1 + h); e += NT) Rh[(e % s0) + (e / s0) * nrm] = 0.0;
    __syncthreads();
    for (int e = tid; e < n * n; e += NT) Km[(e % n) + (e / n) * ld] = v0.Pxx[e];
    for (int e = tid; e < c0 * n; e += NT) Km[(n + e % c0) + (e / c0) * ld] = __ldg(G0 + e);
    for (int i = tid; i < c0; i += NT) Km[(n + i) + (n + i) * ld] = -mudyn;
    for (int i = tid; i < n; i += NT) rhs(i, 0) = -v0.px[i];
    for (int i = tid; i < c0; i += NT) rhs(n + i, 0) = -__ldg(g0 + i);
    for (int e = tid; e < n * h; e += NT) rhs(e % n, 1 + e / n) = -v0.Pxt[e];
    __syncthreads();
    if (!bk_factor_cta<NT>(tid, s0, Km, ld, piv, sub, wk, ctl)) flags |= GARB200_FLAG_BK_SINGULAR;
    __syncthreads();
    for (int col = tid; col < 1 + h; col += NT) bk_solve_column(s0, Km, ld, piv, sub, Rh + (size_t)col * nrm, 1);
    __syncthreads();
    for (int i = tid; i < s0; i += NT) p.kkt0_ff[(size_t)b * s0 + i] = rhs(i, 0);
    for (int e = tid; e < s0 * h; e += NT) p.kkt0_fth[(size_t)b * s0 * h + e] = rhs(e / h, 1 + e % h);   // row-major
    for (int e = tid; e < h * h + h; e += NT) {
      if (e < h * h) {                                  // thHess = Ptt0 + Pxt0^T fth_x      (:92)
        const int a = e % h, bb = e / h;
        double acc = v0.Ptt[e];
        for (int i = 0; i < n; i++) acc = fma(v0.Pxt[i + a * n], rhs(i, 1 + bb), acc);
        p.thHess[(size_t)b * h * h + e] = acc;
      } else {                                          // thGrad = pt0 + Pxt0^T x0ff         (:91)
        const int a = e - h * h;
        double acc = v0.pt[a];
        for (int i = 0; i < n; i++) acc = fma(v0.Pxt[i + a * n], rhs(i, 0), acc);
        p.thGrad[(size_t)b * h + a] = acc;
      }
    }
    for (int e = tid; e < n * n; e += NT)
      if (!isfinite(v0.Pxx[e])) flags |= GARB200_FLAG_NONFINITE;
  }
  if (flags) atomicOr(reinterpret_cast<int *>(p.flags) + b, flags);
}

} // namespace

size_t dense_smem_bytes(const DevParams &p) { return (size_t)dense_plan(p).total * sizeof(double); }

int launch_backward_dense(const DevParams &p, cudaStream_t st) {
  constexpr int NT = 128;
  const size_t smem = dense_smem_bytes(p);
  if (smem > 227u * 1024u) return -1;
  auto kern = dense_backward_kernel<NT>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -2;
  kern<<<p.batch, NT, smem, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

} // namespace garb200
