// condensed.cu — reduced boundary system of the parallel-in-time solver (sm_100a).
//
// After the leg kernel (backward.cu, one CTA per leg) has condensed every leg onto its
// boundary unknowns, ONE CTA per problem assembles and solves the symmetric
// block-tridiagonal saddle system in the unknowns [lbd_0 (nc0), x_0, th_1, x_{i1}, th_2, ...]:
//   assemble      ParallelRiccatiSolver::assembleCondensedSystem   parallel-solver.hxx:63-109
//   solve         symmetricBlockTridiagSolve (up-looking LDL^T)    block-tridiagonal.hpp:80-137
//   residual      blockTridiagMatMul                               block-tridiagonal.hpp:50-75
//   refinement    blockTridiagRefinementStep, <= 5 steps, 1e-11    block-tridiagonal.hpp:144-179,
//                                                                  parallel-solver.hxx:155-175
//   collapse      ParallelRiccatiSolver::collapseFeedback          parallel-solver.hpp:55-66
// The system is never materialised: diagonal / super-diagonal blocks are read where the
// leg kernel left them (value records, knot records, G0); only the factors
// (Bunch-Kaufman of every Schur-complemented diagonal block and the U^T blocks) are
// written, because the refinement steps and the forward substitution re-read them.
#include <cstdlib>
#include "fast.cuh"
#include "kernels.cuh"

namespace garb200 {

namespace {

struct Csys {
  const DevParams &p;
  int b, J, nb, n, c0;
  __device__ Csys(const DevParams &p_, int b_) : p(p_), b(b_), J(p_.num_legs), nb(2 * p_.num_legs), n(p_.nx), c0(p_.nc0) {}
  __device__ int dim(int k) const { return k == 0 ? c0 : n; }
  __device__ size_t voff(int k) const { return k == 0 ? 0 : (size_t)c0 + (size_t)(k - 1) * n; }
  __device__ size_t vlen() const { return (size_t)c0 + (size_t)(nb - 1) * n; }
  __device__ const double *slot(int leg) const {
    int beg, end; get_work(p.T, leg, J, beg, end);
    const int s = p.store_all_values ? beg : leg;
    return p.values + ((size_t)b * p.nslots + s) * p.VL.stride;
  }
  // diagonal block k (dim x dim, col-major): -mu I | Pmat | Vtt        (:72-88)
  template <class G> __device__ void load_diag(const G &g, int k, double *dst) const {
    const int d = dim(k);
    if (k == 0) {
      for (int e = g.tid; e < d * d; e += G::NT) dst[e] = (e % d == e / d) ? -p.mudyn : 0.0;
      return;
    }
    const double *src = (k & 1) ? slot((k - 1) / 2) + p.VL.Pmat : slot(k / 2 - 1) + p.VL.Vtt;
    for (int e = g.tid; e < d * d; e += G::NT) dst[e] = src[e];
  }
  // pointer to super-diagonal block k (dim(k) x dim(k+1), col-major, ld = dim(k))  (:74-91)
  __device__ const double *super_ptr(int k) const {
    if (k == 0) return p.G0 + (size_t)b * c0 * n;
    if (k & 1) return slot((k - 1) / 2) + p.VL.Vxt;
    int i0, i1; get_work(p.T, k / 2 - 1, J, i0, i1);
    return p.knots + ((size_t)b * (p.T + 1) + phys_t(p, i1 - 1)) * p.KL.stride + p.KL.E;
  }
  __device__ const double *diag_ptr(int k) const {   // k >= 1
    return (k & 1) ? slot((k - 1) / 2) + p.VL.Pmat : slot(k / 2 - 1) + p.VL.Vtt;
  }
  __device__ double rhs(int k, int r) const {          // :98-108
    if (k == 0) return -p.g0[(size_t)b * c0 + r];
    if (k & 1) return -(slot((k - 1) / 2) + p.VL.pvec)[r];
    return -(slot(k / 2 - 1) + p.VL.vt)[r];
  }
};

// y[0:M] -= A (M x K, element (i,j) at A[i*ars + j*acs]) * x[0:K]; 4 threads per row.
template <class G>
__device__ __forceinline__ void gemv_sub4(const G &g, int M, int K, const double *A, int ars, int acs,
                                          const double *x, double *y) {
  const int q = g.tid & 3, rl = g.tid >> 2;
  for (int r0 = 0; r0 < M; r0 += G::NT / 4) {
    const int r = r0 + rl;
    double acc = 0.0;
    if (r < M)
      for (int j = q; j < K; j += 4) acc = fma(A[(size_t)r * ars + (size_t)j * acs], x[j], acc);
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (r < M && q == 0) y[r] -= acc;
  }
}

template <int NT>
__global__ void __launch_bounds__(NT) condensed_kernel(const DevParams p, const int only_marked) {
  extern __shared__ __align__(16) double sm[];
  using G = Group<NT>;
  G g; g.tid = threadIdx.x; g.bar_id = 0;
  const int b = blockIdx.x;
  // second launch behind condensed_fast_kernel: only the problems it handed back (refine_steps == -1)
  if (only_marked && p.refine_steps[b] != -1) return;
  const Csys cs(p, b);
  const int n = cs.n, nb = cs.nb, N = nb - 1;
  double *sD = sm, *sD2 = sD + n * n, *sB = sD2 + n * n, *sC = sB + n * n;
  double *ssub = sC + n * n;                       // n + 2
  double *sx = ssub + ((n + 3) & ~1);              // n  (vector being solved)
  double *sy = sx + ((n + 1) & ~1);                // n
  double *sred = sy + ((n + 1) & ~1);              // NT/32
  int *spiv = reinterpret_cast<int *>(sred + NT / 32 + 2);
  __shared__ int s_ok;
  const size_t vlen = cs.vlen();
  double *sol = p.csol + (size_t)b * vlen, *err = p.cerr + (size_t)b * vlen, *rhs = p.crhs + (size_t)b * vlen;
  const size_t bb = (size_t)n * n;
  double *facD = p.cdiagF + (size_t)b * nb * bb, *facU = p.csubF + (size_t)b * nb * bb;
  double *facS = p.csubd + (size_t)b * nb * n;
  int *facP = p.cpiv + (size_t)b * nb * n;
  if (g.tid == 0) s_ok = 1;

  for (int k = 0; k < nb; k++)
    for (int r = g.tid; r < cs.dim(k); r += NT) { const double v = cs.rhs(k, r); rhs[cs.voff(k) + r] = v; sol[cs.voff(k) + r] = v; }
  cs.load_diag(g, N, sD);
  g.sync();

  // factor block k (held in sD, dimension d), keep the factor, solve sol[k] in place
  auto factor_and_store = [&](int k, int d) {
    if (g.warp() == 0) {
      const bool ok = bk_factor_warp(g.lane(), d, sD, d, spiv, ssub);
      if (!ok && g.lane() == 0) s_ok = 0;
    }
    g.sync();
    for (int e = g.tid; e < d * d; e += NT) facD[k * bb + e] = sD[e];
    for (int i = g.tid; i < d; i += NT) { facS[k * n + i] = ssub[i]; facP[k * n + i] = spiv[i]; }
  };

  for (int i = N - 1; i >= 0; i--) {                  // block-tridiagonal.hpp:98-123
    const int di = cs.dim(i), dj = n;
    factor_and_store(i + 1, dj);
    const double *Bg = cs.super_ptr(i);
    for (int e = g.tid; e < di * dj; e += NT) {
      const int r = e % di, c = e / di;
      const double v = Bg[e];
      sB[e] = v;            // B   di x dj
      sC[c + r * dj] = v;   // B^T dj x di
    }
    for (int r = g.tid; r < dj; r += NT) sx[r] = sol[cs.voff(i + 1) + r];
    g.sync();
    if (!s_ok) break;
    for (int col = g.tid; col < di + 1; col += NT) {
      if (col < di) bk_solve_column(dj, sD, dj, spiv, ssub, sC + (size_t)col * dj, 1);
      else bk_solve_column(dj, sD, dj, spiv, ssub, sx, 1);
    }
    g.sync();
    for (int r = g.tid; r < dj; r += NT) sol[cs.voff(i + 1) + r] = sx[r];
    for (int e = g.tid; e < dj * di; e += NT) facU[i * bb + e] = sC[e];
    cs.load_diag(g, i, sD2);
    g.sync();
    gemv_sub4(g, di, dj, sB, 1, di, sx, sol + cs.voff(i));       // rhs[i] -= B rhs[i+1]
    gemm<4, 3>(g, di, di, dj, -1.0, sB, 1, di, sC, 1, dj, 1.0, sD2, 1, di, sD, 1, di);  // diag[i] -= B C
    g.sync();
  }
  if (s_ok) {
    const int d0 = cs.dim(0);
    factor_and_store(0, d0);
    for (int r = g.tid; r < d0; r += NT) sx[r] = sol[r];
    g.sync();
    if (s_ok && g.tid == 0) bk_solve_column(d0, sD, d0, spiv, ssub, sx, 1);
    g.sync();
    for (int r = g.tid; r < d0; r += NT) sol[r] = sx[r];
    g.sync();
  }
  if (!s_ok) {
    if (g.tid == 0) atomicOr(reinterpret_cast<int *>(p.flags) + b, GARB200_FLAG_TRIDIAG_FAIL);
    return;
  }
  for (int i = 0; i < N; i++) {                       // :131-134  rhs[i+1] -= U^T_{i+1} rhs[i]
    gemv_sub4(g, n, cs.dim(i), facU + i * bb, 1, n, sol + cs.voff(i), sol + cs.voff(i + 1));
    g.sync();
  }

  // ---- iterative refinement (parallel-solver.hxx:155-175) -------------------------------
  int steps = 0;
  for (int it = 0; it < 5; it++) {
    // err = rhs - A sol ; resdl = ||err||_inf
    double mx = 0.0;
    for (int k = 0; k < nb; k++) {
      const int d = cs.dim(k);
      const double *Dk = k ? cs.diag_ptr(k) : nullptr;
      const double *Sk = k < N ? cs.super_ptr(k) : nullptr;       // d x dim(k+1)
      const double *Sp = k ? cs.super_ptr(k - 1) : nullptr;       // dim(k-1) x d  (sub = its transpose)
      const int dp = k ? cs.dim(k - 1) : 0;
      for (int r = g.tid; r < d; r += NT) {
        double acc = 0.0;
        if (k == 0) acc = -p.mudyn * sol[r];
        else for (int j = 0; j < d; j++) acc = fma(Dk[r + (size_t)j * d], sol[cs.voff(k) + j], acc);
        if (Sk) for (int j = 0; j < n; j++) acc = fma(Sk[r + (size_t)j * d], sol[cs.voff(k + 1) + j], acc);
        if (Sp) for (int j = 0; j < dp; j++) acc = fma(Sp[j + (size_t)r * dp], sol[cs.voff(k - 1) + j], acc);
        const double e = rhs[cs.voff(k) + r] - acc;
        err[cs.voff(k) + r] = e;
        const double a = fabs(e);
        if (!(a <= mx)) mx = a;
      }
    }
    mx = warp_max(mx);
    if (g.lane() == 0) sred[g.warp()] = mx;
    g.sync();
    double resdl = 0.0;
    for (int w = 0; w < NT / 32; w++) resdl = fmax(resdl, sred[w]);
    g.sync();
    if (resdl <= 1e-11) break;                        // condensedThreshold (parallel-solver.hpp:101)
    // refinement step on err (block-tridiagonal.hpp:144-179)
    for (int i = N; i >= 0; i--) {
      const int d = cs.dim(i);
      for (int e = g.tid; e < d * d; e += NT) sD[e] = facD[i * bb + e];
      for (int r = g.tid; r < d; r += NT) { ssub[r] = facS[i * n + r]; spiv[r] = facP[i * n + r]; sx[r] = err[cs.voff(i) + r]; }
      g.sync();
      if (g.tid == 0) bk_solve_column(d, sD, d, spiv, ssub, sx, 1);
      g.sync();
      for (int r = g.tid; r < d; r += NT) err[cs.voff(i) + r] = sx[r];
      if (i > 0) gemv_sub4(g, cs.dim(i - 1), d, cs.super_ptr(i - 1), 1, cs.dim(i - 1), sx, err + cs.voff(i - 1));
      g.sync();
    }
    for (int i = 0; i < N; i++) {
      gemv_sub4(g, n, cs.dim(i), facU + i * bb, 1, n, err + cs.voff(i), err + cs.voff(i + 1));
      g.sync();
    }
    for (size_t e = g.tid; e < vlen; e += NT) sol[e] += err[e];
    g.sync();
    steps++;
  }
  if (g.tid == 0) p.refine_steps[b] = steps;
}

// ------------------------------------------------------------------------------------------
// Dimension-specialised condensed solve (nx = N, nc0 = N): the same up-looking block elimination,
// refinement and forward substitution (block-tridiagonal.hpp:80-179), restated for the GPU's
// latency budget.  The Schur-complemented diagonal blocks are symmetric INDEFINITE for the
// reference's problem distribution (its generator zeroes a block of Q), so a pivoted factorisation
// is required; here every block is inverted by the block Gauss-Jordan with partial pivoting on the
// tensor pipe (gj_inverse_dmma, one warp, matrix in registers), and
//   * U^T = D^{-1} B^T and the Schur update D_i -= B U^T are DMMA GEMMs, the solved right-hand side a
//     GEMV; D^{-1} is also the block solve of the refinement steps, whose residuals are evaluated
//     against the ORIGINAL blocks, so the rounding of the explicit inverse is corrected there (the
//     reference runs the same <= 5 refinement steps on its Bunch-Kaufman factors);
//   * the next blocks arrive by TMA while the current one is inverted;
//   * solution, residual and right-hand side live in shared memory; the residual is evaluated by
//     all threads over (block, row) pairs at once; the sequential GEMV chains fetch the next matrix
//     rows into registers before the barrier of the current step.
// A zero pivot hands the problem to the Bunch-Kaufman kernel above (refine_steps = -1).
// ------------------------------------------------------------------------------------------
template <int N, int NW>
__global__ void __launch_bounds__(32 * NW, 1) condensed_fast_kernel(const DevParams p) {
  extern __shared__ __align__(16) double sm[];
  using GJ = GjDmma<N>;
  constexpr int n = N, nn = N * N, NT = 32 * NW, RPT = (N + 3) / 4;
  static_assert(N % 2 == 0, "16-byte aligned TMA destinations");
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x;
  const Csys cs(p, b);
  const int nb = cs.nb, Nn = nb - 1;
  const int vlen = nb * n;
  double *sD = sm, *sInv = sD + nn, *sC = sInv + nn, *sB0 = sC + nn, *sG0 = sB0 + 2 * nn;
  double *gjw = sG0 + 2 * nn, *sx = gjw + ((GJ::SMEM + 1) & ~1), *sy = sx + n, *red = sy + n;
  double *ssol = red + 16, *serr = ssol + vlen, *srhs = serr + vlen;
  uint64_t *bars = reinterpret_cast<uint64_t *>(srhs + vlen);
  int *gji = reinterpret_cast<int *>(bars + 2);
  __shared__ int s_fail;
  const size_t bb = (size_t)nn;
  double *facD = p.cdiagF + (size_t)b * nb * bb, *facU = p.csubF + (size_t)b * nb * bb;
  const double mu = p.mudyn;

  if (tid == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); fence_mbar_init(); s_fail = 0; }
  for (int e = tid; e < vlen; e += NT) { const double v = cs.rhs(e / n, e % n); srhs[e] = v; ssol[e] = v; }
  for (int e = tid; e < nn; e += NT) sD[e] = cs.diag_ptr(Nn)[e];
  __syncthreads();
  auto prefetch = [&](int i) {        // thread 0: super(i) and diag(i) -> buffers i & 1
    const int bf = i & 1;
    fence_proxy_async();
    mbar_expect_tx(bars + bf, (uint32_t)(nn * 8) * (i > 0 ? 2u : 1u));
    tma_load_1d(sB0 + bf * nn, cs.super_ptr(i), (uint32_t)nn * 8u, bars + bf);
    if (i > 0) tma_load_1d(sG0 + bf * nn, cs.diag_ptr(i), (uint32_t)nn * 8u, bars + bf);
  };
  if (tid == 0) prefetch(Nn - 1);
  bool failed = false;
#pragma unroll 1
  for (int i = Nn - 1; i >= 0; i--) {                  // block-tridiagonal.hpp:98-123
    const int bf = i & 1;
    const double *sB = sB0 + bf * nn, *sG = sG0 + bf * nn;
    if (warp == 0) {
      if (!gj_inverse_dmma<n>(lane, sD, sInv, true, gjw, gji) && lane == 0) s_fail = 1;
    }
    mbar_wait(bars + bf, (uint32_t)(((Nn - 1 - i) >> 1) & 1));
    if (tid == 0 && i > 0) prefetch(i - 1);
    __syncthreads();
    if (s_fail) {
      if (i > 0) mbar_wait(bars + (bf ^ 1), (uint32_t)(((Nn - i) >> 1) & 1));   // drain the pending prefetch
      failed = true;
      break;
    }
    // U^T = D^{-1} B^T ; rhs[i+1] <- D^{-1} rhs[i+1]
    mma_gemm<n, n, n, NW>(warp, lane,
        [&](int r, int k) { return sInv[r + k * n]; },
        [&](int k, int c) { return sB[c + k * n]; },
        [](int, int) { return 0.0; },
        [&](int r, int c, double v) { sC[r + c * n] = v; });
    gemv4<n, n, NT>(tid, [&](int r, int j) { return sInv[r + j * n]; }, [&](int j) { return ssol[(i + 1) * n + j]; },
                    [](int) { return 0.0; }, [&](int r, double v) { sx[r] = v; });
    __syncthreads();
    for (int e = tid; e < nn; e += NT) { facU[(size_t)i * bb + e] = sC[e]; facD[(size_t)(i + 1) * bb + e] = sInv[e]; }
    // diag[i] -= B U^T ; rhs[i] -= B rhs[i+1]
    mma_gemm<n, n, n, NW>(warp, lane,
        [&](int r, int k) { return -sB[r + k * n]; },
        [&](int k, int c) { return sC[k + c * n]; },
        [&](int r, int c) { return i == 0 ? (r == c ? -mu : 0.0) : sG[r + c * n]; },
        [&](int r, int c, double v) { sD[r + c * n] = v; });
    gemv4<n, n, NT>(tid, [&](int r, int j) { return -sB[r + j * n]; }, [&](int j) { return sx[j]; },
                    [&](int r) { return ssol[i * n + r]; }, [&](int r, double v) { sy[r] = v; });
    __syncthreads();
    if (tid < n) { ssol[(i + 1) * n + tid] = sx[tid]; ssol[i * n + tid] = sy[tid]; }
  }
  if (!failed) {
    __syncthreads();
    if (warp == 0) {
      if (!gj_inverse_dmma<n>(lane, sD, sInv, true, gjw, gji) && lane == 0) s_fail = 1;
    }
    __syncthreads();
    if (s_fail) failed = true;
  }
  if (failed) {                        // uniform over the CTA: hand the problem to the Bunch-Kaufman kernel
    if (tid == 0) p.refine_steps[b] = -1;
    return;
  }
  gemv4<n, n, NT>(tid, [&](int r, int j) { return sInv[r + j * n]; }, [&](int j) { return ssol[j]; },
                  [](int) { return 0.0; }, [&](int r, double v) { sx[r] = v; });
  for (int e = tid; e < nn; e += NT) facD[e] = sInv[e];
  __syncthreads();
  if (tid < n) ssol[tid] = sx[tid];
  __syncthreads();

  // sequential GEMV chain helper: y_dst[r] = y_src[r] - sum_j A(r, j) x[j], 4 threads per row, the
  // matrix rows already in registers (fetched before the previous barrier)
  const int gr = tid >> 2, gq = tid & 3;
  auto fetch = [&](const double *A, double (&a)[RPT]) {         // A col-major n x n, global
#pragma unroll
    for (int u = 0; u < RPT; u++) { const int j = gq + 4 * u; a[u] = (gr < n && j < n) ? __ldg(A + gr + (size_t)j * n) : 0.0; }
  };
  auto apply = [&](const double (&a)[RPT], const double *x, const double *ysrc, double *ydst, double sgn) {
    double acc = 0.0;
#pragma unroll
    for (int u = 0; u < RPT; u++) { const int j = gq + 4 * u; if (j < n) acc = fma(a[u], x[j], acc); }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (gr < n && gq == 0) ydst[gr] = (ysrc ? ysrc[gr] : 0.0) + sgn * acc;
  };
  static_assert(NT >= 4 * N, "GEMV chain: 4 threads per row");
  {
    double a[RPT];
    fetch(facU, a);
    for (int i = 0; i < Nn; i++) {                     // :131-134  rhs[i+1] -= U^T_{i+1} rhs[i]
      double an[RPT];
      if (i + 1 < Nn) fetch(facU + (size_t)(i + 1) * bb, an);
      apply(a, ssol + i * n, ssol + (i + 1) * n, ssol + (i + 1) * n, -1.0);
      __syncthreads();
#pragma unroll
      for (int u = 0; u < RPT; u++) a[u] = an[u];
    }
  }

  // ---- iterative refinement (parallel-solver.hxx:155-175) -------------------------------
  int steps = 0;
  for (int it = 0; it < 5; it++) {
    double mx = 0.0;
    for (int idx = tid; idx < vlen; idx += NT) {       // err = rhs - A sol, all (block, row) pairs at once
      const int k = idx / n, r = idx % n;
      double acc = 0.0;
      if (k == 0) acc = -mu * ssol[r];
      else {
        const double *Dk = cs.diag_ptr(k);
        for (int j = 0; j < n; j++) acc = fma(__ldg(Dk + r + (size_t)j * n), ssol[k * n + j], acc);
      }
      if (k < Nn) {
        const double *Sk = cs.super_ptr(k);
        for (int j = 0; j < n; j++) acc = fma(__ldg(Sk + r + (size_t)j * n), ssol[(k + 1) * n + j], acc);
      }
      if (k > 0) {
        const double *Sp = cs.super_ptr(k - 1);
        for (int j = 0; j < n; j++) acc = fma(__ldg(Sp + j + (size_t)r * n), ssol[(k - 1) * n + j], acc);
      }
      const double e = srhs[idx] - acc;
      serr[idx] = e;
      const double ae = fabs(e);
      if (!(ae <= mx)) mx = ae;
    }
    mx = warp_max(mx);
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    double resdl = 0.0;
    for (int w = 0; w < NW; w++) resdl = fmax(resdl, red[w]);
    __syncthreads();
    if (resdl <= 1e-11) break;                        // condensedThreshold (parallel-solver.hpp:101)
    // refinement step on err (block-tridiagonal.hpp:144-179) with D^{-1} as the block solve
    {
      double a[RPT], a2[RPT];
      fetch(facD + (size_t)Nn * bb, a);
      for (int i = Nn; i >= 0; i--) {
        if (i > 0) fetch(cs.super_ptr(i - 1), a2);
        apply(a, serr + i * n, nullptr, sx, 1.0);                    // z = D_i^{-1} e_i
        __syncthreads();
        if (tid < n) serr[i * n + tid] = sx[tid];
        if (i > 0) {
          fetch(facD + (size_t)(i - 1) * bb, a);
          apply(a2, sx, serr + (i - 1) * n, serr + (i - 1) * n, -1.0);   // e_{i-1} -= B_{i-1} z
        }
        __syncthreads();
      }
    }
    {
      double a[RPT];
      fetch(facU, a);
      for (int i = 0; i < Nn; i++) {
        double an[RPT];
        if (i + 1 < Nn) fetch(facU + (size_t)(i + 1) * bb, an);
        apply(a, serr + i * n, serr + (i + 1) * n, serr + (i + 1) * n, -1.0);
        __syncthreads();
#pragma unroll
        for (int u = 0; u < RPT; u++) a[u] = an[u];
      }
    }
    for (int e = tid; e < vlen; e += NT) ssol[e] += serr[e];
    __syncthreads();
    steps++;
  }
  double *sol = p.csol + (size_t)b * cs.vlen();
  for (int e = tid; e < vlen; e += NT) sol[e] = ssol[e];
  if (tid == 0) p.refine_steps[b] = steps;
}

// K0 -= Kth0 * sub[1]; after the reference's swap (parallel-solver.hxx:152-153)
// condensedKktSystem.subdiagonal[1] is the ORIGINAL block super[1]^T = Vxt(knot 0)^T.
__global__ void collapse_kernel(const DevParams p) {
  const int b = blockIdx.x, n = p.nx, h = p.GL.nth;
  const Csys cs(p, b);
  const double *Vxt = cs.slot(0) + p.VL.Vxt;            // n x h col-major
  double *rec = p.gains + ((size_t)b * (p.T + 1) + phys_t(p, 0)) * p.GL.stride;
  double *fb = rec + p.GL.fb;
  const double *fth = rec + p.GL.fth;
  const int m = p.nu_t[0];
  for (int e = threadIdx.x; e < m * n; e += blockDim.x) {
    const int i = e / n, j = e % n;
    double acc = 0.0;
    for (int l = 0; l < h; l++) acc = fma(fth[i * h + l], Vxt[j + (size_t)l * n], acc);  // sub[1](l, j) = Vxt(j, l)
    fb[e] -= acc;
  }
}

} // namespace

int launch_condensed(const DevParams &p, cudaStream_t st) {
  constexpr int NT = 128;
  const int n = p.nx;
  const size_t d = (size_t)4 * n * n + 3 * (n + 4) + NT / 32 + 4 + (n + 4) / 2 + 2;
  const size_t smem = d * sizeof(double);
  if (smem > 227u * 1024u) return -1;
  auto kern = condensed_kernel<NT>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return -2;
  int launched = 0, only_marked = 0;
  const bool force_bk = p.tune.condensed == 1;
  if (n == 36 && p.nc0 == 36 && !force_bk) {
    constexpr int N = 36, NW = 5;
    const size_t vlen = (size_t)2 * p.num_legs * N;
    const size_t fd = (size_t)7 * N * N + GjDmma<N>::SMEM + 2 + 2 * N + 16 + 3 * vlen + 4 + GjDmma<N>::SINTS / 2 + 2;
    if (fd * sizeof(double) <= 200u * 1024u) {
      auto kf = condensed_fast_kernel<N, NW>;
      if (cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(fd * sizeof(double))) != cudaSuccess) return -2;
      kf<<<p.batch, 32 * NW, fd * sizeof(double), st>>>(p);
      if (cudaGetLastError() != cudaSuccess) return -2;
      launched++;
      only_marked = 1;
    }
  }
  const bool fast_only = p.tune.condensed == 2;   // debugging
  if (fast_only && only_marked) return launched;
  kern<<<p.batch, NT, smem, st>>>(p, only_marked);
  return cudaGetLastError() == cudaSuccess ? launched + 1 : -2;
}

int launch_collapse(const DevParams &p, cudaStream_t st) {
  collapse_kernel<<<p.batch, 128, 0, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

} // namespace garb200
