// sweep_big.cu — specialised proximal Riccati backward sweep for knots whose record does not
// fit in shared memory next to the work matrices (nx = 76, nu = 32: the Talos-walk shape of
// BASELINE config 5; knot record 187 KB, gains record 113 KB).
//
// Same algorithm and step numbering as sweep_fast.cu (riccati-kernel.hxx:90-150, 167-250;
// proximal-riccati.hxx:44-62) with a different memory plan:
//   * the big INPUT blocks of the knot ([Q|S], [A|B], E^{-1}) stay in global memory (L2: the whole
//     record is prefetched one knot ahead with cp.async.bulk.prefetch.L2) and enter the DMMA GEMMs
//     as the STRIP-side operand, so every element is fetched once per product, straight into the
//     fragment registers of the one warp that needs it (strip_gemm, fast.cuh);
//   * only the work matrices live in shared memory: X1, P, [Qhat|Shat], BtV, the KKT block and the
//     staged part of the gains record (221 KB, one CTA of 10 warps per SM);
//   * products whose big input would sit on the broadcast side are transposed
//     (Ptilde^T = Einv^T X1^T, [Qhat|Shat]^T = [Q|S]^T + [A|B]^T (Vxx A)) — the same numbers, the
//     strips just run over the other index.
#include <cstdlib>
#include "fast.cuh"
#include "kernels.cuh"

namespace garb200 {

// ---- pre-pass: Einv = PartialPivLU(E).solve(I), one CTA per (problem, knot) -------------------
template <int NT>
__global__ void __launch_bounds__(NT) einv_cta_kernel(const DevParams p) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.nx, nn = n * n, tid = threadIdx.x;
  double *E = sm, *Y = E + nn, *X = Y + nn;
  int *iw = reinterpret_cast<int *>(X + nn);
  const long long w = blockIdx.x;
  const int b = (int)(w / p.T), t = (int)(w % p.T);
  const double *Eg = p.knots + ((size_t)b * (p.T + 1) + phys_t(p, t)) * p.KL.stride + p.KL.E;
  double *Xg = p.einv + ((size_t)b * p.T + phys_t(p, t)) * (size_t)nn;
  for (int e = tid; e < nn; e += NT) E[e] = Eg[e];
  __syncthreads();
  Group<NT> g{tid, 0};
  const bool ok = lu_inverse(g, n, E, Y, X, iw);
  __syncthreads();
  for (int e = tid; e < nn; e += NT) Xg[e] = X[e];
  if (!ok && tid == 0) atomicOr(reinterpret_cast<int *>(p.flags) + b, GARB200_FLAG_LU_SINGULAR);
}

// The same pre-pass as a block Gauss-Jordan on the tensor pipe (gj_inverse_cta, fast.cuh).
template <int N, int NWARP>
__global__ void __launch_bounds__(32 * NWARP, 3) einv_cta_dmma_kernel(const DevParams p) {
  using G = GjCta<N>;
  extern __shared__ __align__(16) double sm[];
  int *si = reinterpret_cast<int *>(sm + G::SMEM);
  const long long w = blockIdx.x;
  const int b = (int)(w / p.T), t = (int)(w % p.T);
  const double *Eg = p.knots + ((size_t)b * (p.T + 1) + phys_t(p, t)) * p.KL.stride + p.KL.E;
  double *Xg = p.einv + ((size_t)b * p.T + phys_t(p, t)) * (size_t)(N * N);
  const bool ok = gj_inverse_cta<N, NWARP>(threadIdx.x, Eg, Xg, sm, si);
  if (!ok && threadIdx.x == 0) atomicOr(reinterpret_cast<int *>(p.flags) + b, GARB200_FLAG_LU_SINGULAR);
}

// The same with the tiles in registers (gj_inverse_cta_regs): half the shared-memory traffic per matrix.
template <int N, int NWARP>
__global__ void __launch_bounds__(32 * NWARP, 3) einv_cta_regs_kernel(const DevParams p) {
  using G = GjCtaRegs<N, NWARP>;
  extern __shared__ __align__(16) double sm[];
  int *si = reinterpret_cast<int *>(sm + G::SMEM);
  const long long w = blockIdx.x;
  const int b = (int)(w / p.T), t = (int)(w % p.T);
  const double *Eg = p.knots + ((size_t)b * (p.T + 1) + phys_t(p, t)) * p.KL.stride + p.KL.E;
  double *Xg = p.einv + ((size_t)b * p.T + phys_t(p, t)) * (size_t)(N * N);
  const bool ok = gj_inverse_cta_regs<N, NWARP>(threadIdx.x, Eg, Xg, sm, si);
  if (!ok && threadIdx.x == 0) atomicOr(reinterpret_cast<int *>(p.flags) + b, GARB200_FLAG_LU_SINGULAR);
}

// ---- record offsets (identical to garb200_make_knot_layout / garb200_make_gain_layout) ---------
template <int NX, int NU> struct BigPlan {
  static constexpr int n = NX, m = NU, s = NU, na = NX + 1;
  static constexpr int sp = (s + 3) & ~3;
  static constexpr int r2(int x) { return (x + 1) & ~1; }
  static constexpr int oQ = 0, oS = r2(oQ + n * n), oR = r2(oS + n * m), oq = r2(oR + m * m),
                       orr = r2(oq + n), oA = r2(orr + m), oB = r2(oA + n * n), oE = r2(oB + n * m),
                       of = r2(oE + n * n), kstride = r2(of + n);
  static_assert(oS == oQ + n * n && oB == oA + n * n, "[Q|S] and [A|B] must be contiguous");
  static constexpr int nrows = m + 2 * n;
  static constexpr int gff = 0, gfb = r2(nrows), gfth = r2(gfb + nrows * n), gstride = gfth;
  static constexpr int gpart = r2(gfb + s * n);        // [ff | fb rows K] staged in shared memory
  // shared memory (doubles).  The first five blocks are overlaid by the (nx + nc0)^2 initial-stage
  // KKT matrix when G0 != I; everything from G on must survive that.
  static constexpr int X1 = 0;                          // n x na
  static constexpr int QS = r2(X1 + n * na);            // n x (n + m): [Qhat | Shat]
  static constexpr int BtV = r2(QS + n * (n + m));      // m x na
  static constexpr int kkt = r2(BtV + m * na);          // sp x sp
  static constexpr int Rqr = r2(kkt + sp * sp);         // [R | q | r] of the knot record
  static constexpr int fv = r2(Rqr + (oA - oR));        // f
  static constexpr int P = r2(fv + n);                  // n x na
  static constexpr int K0max = (2 * n) * ((2 * n) | 1);  // overlay extent of the initial-stage KKT matrix
  static constexpr int G = r2(P + n * na) > K0max ? r2(P + n * na) : r2(K0max);   // gains staging, behind the overlay
  static constexpr int KS = 2 * n + 8;
  static constexpr int ksub = r2(G + gpart);
  static constexpr int dinv = r2(ksub + KS);
  static constexpr int red = r2(dinv + n);
  static constexpr int atvx = r2(red + 16);
  static constexpr int ints = r2(atvx + n + m);
  static constexpr int bar = r2(ints + (KS + 1) / 2);
  static constexpr int total = bar + 2;
  static_assert(K0max <= G, "initial-stage KKT overlay must end before the gains staging block");
  static_assert(n * ((2 * n) | 1) <= P, "columns < n of the overlay must end before the P block");
  static_assert(gpart >= r2(2 * n) + 4 * n + 4, "gains staging doubles as BK scratch");
  static_assert(total * 8 <= 227 * 1024, "shared-memory plan exceeds 227 KB");
};

// EX: explicit dynamics, E_t == -I for every knot (explicit-dynamics.hxx:32; checked on the device or promised by
// the caller): no pre-pass, E^{-1} is never read, Ptilde = Pmat, -E^{-1} [A_pre | tv] = [A_pre | tv] (see sweep_fast.cu).
template <int NX, int NU, int NW, int MINB, bool PROF = false, bool EX = false>
__global__ void __launch_bounds__(32 * NW, MINB) sweep_big_kernel(const DevParams p) {
  extern __shared__ __align__(16) double sm[];
  using PL = BigPlan<NX, NU>;
  constexpr int n = NX, m = NU, s = NU, NT = 32 * NW, na = PL::na, sp = PL::sp;
  constexpr int rK = 0, rL = m, rA = m + n;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x, T = p.T;
  const double mudyn = p.mudyn, mueq = p.mueq;
  (void)mueq;
  double *X1 = sm + PL::X1, *Qh = sm + PL::QS, *Sh = Qh + n * n, *BtV = sm + PL::BtV, *kkt = sm + PL::kkt;
  double *R = sm + PL::Rqr, *qv = R + (PL::oq - PL::oR), *rv = R + (PL::orr - PL::oR), *f = sm + PL::fv;
  double *P = sm + PL::P, *Gs = sm + PL::G, *ksub = sm + PL::ksub, *dinv = sm + PL::dinv;
  double *red = sm + PL::red, *atvx = sm + PL::atvx;
  double *pv = P + n * n, *x1v = X1 + n * n, *btvx = BtV + m * n;
  double *ff = Gs + PL::gff, *fb = Gs + PL::gfb;
  int *kpiv = reinterpret_cast<int *>(sm + PL::ints);
  uint64_t *bar = reinterpret_cast<uint64_t *>(sm + PL::bar);
  int flags = 0;
  // optional phase timers (GARB200_PHASE_TIMERS=1), same slots as sweep_fast.cu / tools/phase_timers.py;
  // thread 0's view, and the instrumented build spills more: read the shares, not the absolute cycles
  long long ph_acc[16], ph_last = 0;
  if constexpr (PROF) {
#pragma unroll
    for (int i = 0; i < 16; i++) ph_acc[i] = 0;
    ph_last = clock64();
  }
  auto PH = [&](int i) {
    if constexpr (PROF) {
      if (tid == 0) { const long long now = clock64(); ph_acc[i] += now - ph_last; ph_last = now; }
    }
  };

  if (tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  for (int e = tid; e < PL::gpart; e += NT) Gs[e] = 0.0;
  cta_sync<NT>();
  uint32_t phase = 0;
  const size_t pbase = (size_t)b * (T + 1);
  auto knot_ptr = [&](int t) { return p.knots + (pbase + phys_t(p, t)) * PL::kstride; };
  auto einv_ptr = [&](int t) { return p.einv + ((size_t)b * T + phys_t(p, t)) * (size_t)(n * n); };
  auto issue_small = [&](int t) {     // thread 0: [R|q|r] and f of knot t -> shared memory
    const double *src = knot_ptr(t);
    constexpr uint32_t b1 = (uint32_t)(PL::oA - PL::oR) * 8u, b2 = (uint32_t)PL::r2(n) * 8u;
    fence_proxy_async();
    mbar_expect_tx(bar, b1 + b2);
    tma_load_1d(R, src + PL::oR, b1, bar);
    tma_load_1d(f, src + PL::of, b2, bar);
  };
  auto prefetch_knot = [&](int t) {   // thread 0: whole record + E^{-1} of knot t towards L2
    const char *k = reinterpret_cast<const char *>(knot_ptr(t));
    constexpr uint32_t kb = (uint32_t)PL::kstride * 8u, CH = 32768u;
    for (uint32_t o = 0; o < kb; o += CH) l2_prefetch(k + o, (kb - o < CH) ? (kb - o) : CH);
    if (!EX && t < T) {
      const char *e = reinterpret_cast<const char *>(einv_ptr(t));
      constexpr uint32_t eb = (uint32_t)(n * n) * 8u;
      for (uint32_t o = 0; o < eb; o += CH) l2_prefetch(e + o, (eb - o < CH) ? (eb - o) : CH);
    }
  };
  auto store_gains = [&](int t) {     // all threads; Gs complete
    fence_proxy_async();
    cta_sync<NT>();
    if (tid == 0) {
      tma_store_1d(p.gains + (pbase + phys_t(p, t)) * p.GL.stride, Gs, (uint32_t)PL::gpart * 8u);
      tma_store_commit();
    }
  };
  auto store_value = [&](int t) {     // [P | pv] is the value function of knot t
    const int slot = p.store_all_values ? t : (t == 0 ? 0 : -1);
    if (slot < 0) return;
    double *v = p.values + ((size_t)b * p.nslots + slot) * p.VL.stride;
    for (int e = tid; e < n * n; e += NT) v[p.VL.Pmat + e] = P[e];
    for (int i = tid; i < n; i += NT) v[p.VL.pvec + i] = pv[i];
    if (t == 0 && p.store_all_values) {                       // proximal-riccati.hxx:39-42
      for (int e = tid; e < n * n; e += NT) v[p.VL.Vxx + e] = P[e];
      for (int i = tid; i < n; i += NT) v[p.VL.vx + i] = pv[i];
    }
  };

  // ---- terminal knot (nu = nc = 0): Pmat = Q, pvec = q, empty gains  (riccati-kernel.hxx:90-150)
  if (tid == 0) { issue_small(T); if (T > 0) prefetch_knot(T - 1); }
  {
    const double *Qg = knot_ptr(T) + PL::oQ;
    for (int e = tid; e < n * n; e += NT) P[e] = __ldg(Qg + e);
  }
  mbar_wait(bar, phase); phase ^= 1;
  for (int i = tid; i < n; i += NT) pv[i] = qv[i];
  store_gains(T);
  {
    double *gT = p.gains + (pbase + phys_t(p, T)) * p.GL.stride;
    for (int e = PL::gpart + tid; e < PL::gstride; e += NT) gT[e] = 0.0;
  }
  store_value(T);
  fence_proxy_async();
  cta_sync<NT>();

#pragma unroll 1
  for (int t = T - 1; t >= 0; t--) {
    const double *kn = knot_ptr(t);
    const double *gQS = kn + PL::oQ, *gAB = kn + PL::oA, *gB = kn + PL::oB, *gE = EX ? nullptr : einv_ptr(t);
    (void)gE;
    if (tid == 0) {
      issue_small(t);
      if (t > 0) prefetch_knot(t - 1);
      tma_store_wait_read();
    }
    PH(0);
    double fro = 0.0;
    if constexpr (EX) {
      // E^{-1} = -I: Ptilde = Pmat, lower-symmetrised (:179-180, exact); ptilde = pvec
      for (int e = tid; e < n * n; e += NT) {
        const int i = e % n, j = e / n;
        if (i >= j) {
          const double v = P[e];
          if (i > j) P[j + i * n] = v;
          fro = fma(i == j ? v : 2.0 * v, v, fro);
        }
      }
      PH(1);
    } else {
    // (1) [X1 | x1v] = Einv^T [P | pvec]          EinvP, -ptilde      (:175-178)
    strip_gemm<n, na, n, NW, true, false>(warp, lane,
        [&](int i, int k) { return __ldg(gE + k + i * n); },
        [&](int k, int j) { return P[k + j * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) { X1[i + j * n] = v; });
    cta_sync<NT>();
    PH(1);
    // (2) Ptilde^T = Einv^T X1^T; the upper triangle of the transposed product IS the reference's
    //     lower triangle of Ptilde = X1 Einv, mirrored; ||Ptilde||_F^2 on the fly      (:179-180)
    strip_gemm<n, n, n, NW, true, false>(warp, lane,
        [&](int i, int k) { return __ldg(gE + k + i * n); },
        [&](int k, int j) { return X1[j + k * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) {
          if (i <= j) { P[j + i * n] = v; P[i + j * n] = v; fro = fma(i == j ? v : 2.0 * v, v, fro); }
        });
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) fro += __shfl_xor_sync(0xffffffffu, fro, o);
    if (lane == 0) red[warp] = fro;
    mbar_wait(bar, phase); phase ^= 1;      // [R|q|r], f of this knot have landed
    cta_sync<NT>();
    fro = 0.0;
#pragma unroll
    for (int w = 0; w < NW; w++) fro += red[w];
    // order of the Neumann series for Lambda^{-1} (see sweep_fast.cu): truncation x^(order+1) <= 1e-18
    const double xnorm = mudyn * sqrt(fro);
    // (orders 2 and 3 are compiled into the explicit-dynamics kernel only, where Ptilde = Pmat is not scaled
    //  down by |E|^-2: the general kernel sits at the 168-register cap and the extra loop costs it 6 %)
    const int order = xnorm <= 1e-9 ? 1 : !EX ? 0 : xnorm <= 1e-6 ? 2 : xnorm <= 3e-5 ? 3 : 0;
    const bool neumann = order > 0;
    PH(2);
    // (3) w = ptilde + Ptilde f -> pv                                                   (:182-187)
    gemv4<n, n, NT>(tid, [&](int i, int j) { return P[i + j * n]; }, [&](int j) { return f[j]; },
                    [&](int i) { return EX ? pv[i] : -x1v[i]; }, [&](int i, double v) { pv[i] = v; });   // ptilde = -x1v (EX: = pvec)
    if (!neumann) {
      for (int e = tid; e < n * n; e += NT) {
        const int i = e % n, j = e / n;
        if (i >= j) X1[e] = (i == j ? 1.0 : 0.0) + mudyn * P[e];
      }
    }
    cta_sync<NT>();
    PH(3);
    // (4) [Vxx | vx] = Lambda^{-1} [Ptilde | w]                                      (:185-190)
    if (neumann) {
      strip_gemm<n, na, n, NW, false, false>(warp, lane,
          [&](int i, int k) { return -mudyn * P[i + k * n]; },
          [&](int k, int j) { return P[k + j * n]; },
          [&](int i, int j) { return P[i + j * n]; },
          [&](int i, int j, double v) { X1[i + j * n] = v; });
      if constexpr (EX) {
#pragma unroll 1
        for (int it = 1; it < order; it++) {                  // Y <- [Ptilde | w] - mudyn Ptilde Y, in place (column strips)
          cta_sync<NT>();
          mma_gemm_colstrips<n, na, n, NW>(warp, lane,
              [&](int i, int k) { return -mudyn * P[i + k * n]; },
              [&](int k, int j) { return X1[k + j * n]; },
              [&](int i, int j) { return P[i + j * n]; },
              [&](int i, int j, double v) { X1[i + j * n] = v; });
        }
      }
      cta_sync<NT>();
      for (int e = tid; e < n * na; e += NT) {                // back into P, mirrored from the lower triangle
        const int i = e % n, j = e / n;
        P[e] = (j < n && i < j) ? X1[j + i * n] : X1[e];
      }
    } else {
      if (!ldlt_solve_blocked<n, na, NT>(tid, X1, RhsColMajor<n>{P}, Gs, dinv)) flags |= GARB200_FLAG_LLT_NOT_PD;
      mirror_lower_ct<n, NT>(tid, P);
    }
    cta_sync<NT>();
    if (p.store_all_values) {
      double *v = p.values + ((size_t)b * p.nslots + (t + 1)) * p.VL.stride;
      for (int e = tid; e < n * n; e += NT) v[p.VL.Vxx + e] = P[e];
      for (int i = tid; i < n; i += NT) v[p.VL.vx + i] = pv[i];
    }
    PH(5);
    // (5) [AtV ; BtV] = [A | B]^T [Vxx | vx]: rows < n -> X1 (+ A^T vx), rows >= n -> BtV (+ B^T vx)
    strip_gemm<n + m, na, n, NW, true, false>(warp, lane,
        [&](int i, int k) { return __ldg(gAB + k + i * n); },
        [&](int k, int j) { return P[k + j * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) {
          if (i < n) { if (j < n) X1[i + j * n] = v; else atvx[i] = v; }
          else BtV[(i - n) + j * m] = v;
        });
    cta_sync<NT>();
    PH(6);
    // (6a) Rhat = R + BtV B -> kkt (all warps), right-hand side -rhat                       (:199-226)
    auto build_kkt = [&]() {
      mma_gemm<m, m, n, NW>(warp, lane,
          [&](int i, int k) { return BtV[i + k * m]; },
          [&](int k, int j) { return __ldg(gB + k + j * n); },
          [&](int i, int j) { return R[i + j * m]; },
          [&](int i, int j, double v) { kkt[i + j * sp] = v; });
      for (int i = tid; i < m; i += NT) ff[rK + i] = -(rv[i] + btvx[i]);          // -rhat
      if constexpr (sp > s) {
        constexpr int pr = sp - s > 0 ? sp - s : 1;
        for (int e = tid; e < (sp - s) * sp; e += NT) {
          const int i = s + e % pr, j = e / pr;
          kkt[i + j * sp] = (i == j) ? -1.0 : 0.0;
        }
      }
    };
    auto build_rhs_K = [&]() {   // needs Shat complete: fb(k, j) = -Shat(j, k), the same linear index (fallback path)
      for (int e = tid; e < m * n; e += NT) fb[rK * n + e] = -Sh[e];
    };
    build_kkt();
    cta_sync<NT>();
    // (6b) [Qhat | Shat]^T = [Q | S]^T + [A | B]^T (Vxx A),  Vxx A = AtV^T  on warps 1..NW-1   (:195-199)
    //      (-Shat^T goes straight into the K rows of the right-hand side), while
    // (7a) warp 0 factorises Rhat: blocked LDL^T, a dependent FP64 chain that needs nothing else  (:202-208)
    int *kkt_ok = reinterpret_cast<int *>(red + NW);
    if (warp == 0) {
      const bool okf = ldlt_factor_warp<sp, m>(lane, kkt, P, dinv);     // P[0, 4 sp): Vxx is consumed
      if (lane == 0) *kkt_ok = okf ? 1 : 0;
    } else {
      strip_gemm<n + m, n, n, NW - 1, true, true>(warp - 1, lane,
          [&](int i, int k) { return __ldg(gAB + k + i * n); },
          [&](int k, int j) { return X1[j + k * n]; },
          [&](int i, int j) { return __ldg(gQS + j + i * n); },
          [&](int i, int j, double v) {
            Qh[j + i * n] = v;
            if (i >= n) fb[rK * n + j + (i - n) * n] = -v;
          });
      for (int i = tid - 32; i < n; i += NT - 32) qv[i] += atvx[i];                 // qhat
    }
    cta_sync<NT>();
    PH(7);
    // (7b) [K kff] = Rhat^{-1} rhs, one right-hand-side column per thread; Bunch-Kaufman
    //      (bunchkaufman.hpp:20-167, 447-516) when the factorization failed its checks      (:210-228)
    {
      const bool okk = *kkt_ok != 0;
      if (okk) {
        ldlt_solve_dmma<sp, na, NT>(tid, kkt, dinv, RhsGainsTh<n, 0, s>{ff, fb, nullptr});
        cta_sync<NT>();
      } else {
        build_kkt();
        build_rhs_K();
        cta_sync<NT>();
        if (warp == 0) {
          if (!bk_factor_warp(lane, s, kkt, sp, kpiv, ksub)) flags |= GARB200_FLAG_BK_SINGULAR;
        }
        cta_sync<NT>();
        for (int col = tid; col < na; col += NT) {
          if (col == 0) bk_solve_column(s, kkt, sp, kpiv, ksub, ff, 1);
          else bk_solve_column(s, kkt, sp, kpiv, ksub, fb + (col - 1), n);
        }
        cta_sync<NT>();
      }
    }
    PH(9);
    // (8) lff = vx + BtV^T kff ; tv = f + B kff - mudyn lff -> pv slot                  (:231-236)
    for (int i = tid; i < n; i += NT) {
      double l = pv[i], tb = f[i];
#pragma unroll 8
      for (int k = 0; k < m; k++) { l = fma(BtV[k + i * m], ff[rK + k], l); tb = fma(__ldg(gB + i + k * n), ff[rK + k], tb); }
      ff[rL + i] = l;
      pv[i] = tb - mudyn * l;
    }
    // (9) L = AtV^T + BtV^T K -> fb rows rL (straight to HBM)                           (:238-239)
    //     A_pre = A + B K - mudyn L -> P block                                          (:241-242)
    double *gfbg = p.gains + (pbase + phys_t(p, t)) * p.GL.stride + PL::gfb;
    {
      constexpr int TM = (n + 7) / 8, TN = (n + 7) / 8, KSt = (m + 3) / 4;
      const int g = lane >> 2, q = lane & 3;
#pragma unroll 1
      for (int sidx = warp; sidx < TM; sidx += NW) {
        const int ic = sidx * 8 + g, ia = min(ic, n - 1);
        double aBk[KSt];
#pragma unroll
        for (int ks = 0; ks < KSt; ks++) {
          const int k = ks * 4 + q;
          aBk[ks] = ((m % 4 == 0) || k < m) ? __ldg(gB + ia + min(k, m - 1) * n) : 0.0;
        }
        double l0[TN], l1[TN], a0[TN], a1[TN];
        int jbk[TN];
#pragma unroll
        for (int u = 0; u < TN; u++) {
          jbk[u] = min(u * 8 + g, n - 1);
          const int jc = u * 8 + 2 * q, j0 = min(jc, n - 1), j1 = min(jc + 1, n - 1);
          l0[u] = X1[j0 + ia * n]; l1[u] = X1[j1 + ia * n];
          a0[u] = __ldg(gAB + ia + j0 * n); a1[u] = __ldg(gAB + ia + j1 * n);
        }
#pragma unroll
        for (int ks = 0; ks < KSt; ks++) {
          const int k = ks * 4 + q;
          const bool kv = (m % 4 == 0) || (k < m);
          const int kc = min(k, m - 1);
          const double aL = kv ? BtV[kc + ia * m] : 0.0;
          double bvK[TN];
#pragma unroll
          for (int u = 0; u < TN; u++) bvK[u] = kv ? fb[(rK + kc) * n + jbk[u]] : 0.0;
#pragma unroll
          for (int u = 0; u < TN; u++) { dmma_884(l0[u], l1[u], aL, bvK[u]); dmma_884(a0[u], a1[u], aBk[ks], bvK[u]); }
        }
#pragma unroll
        for (int u = 0; u < TN; u++) {
          const int jc = u * 8 + 2 * q;
          if (ic < n && jc < n) {
            *reinterpret_cast<double2 *>(gfbg + (rL + ic) * n + jc) = make_double2(l0[u], l1[u]);
            P[ic + jc * n] = a0[u] - mudyn * l0[u];
            P[ic + (jc + 1) * n] = a1[u] - mudyn * l1[u];
          }
        }
      }
    }
    cta_sync<NT>();
    PH(10);
    // (10) [Afb | yff] = -Einv [A_pre | tv]                                              (:243-245)
    if constexpr (EX) {                 // -E^{-1} = I: the block itself, row-major into the A rows of fb
      for (int e = tid; e < n * n; e += NT) { const int i = e / n, j = e % n; gfbg[(rA + i) * n + j] = P[i + j * n]; }
      for (int i = tid; i < n; i += NT) ff[rA + i] = pv[i];
    } else {
    strip_gemm<n, na, n, NW, true, false>(warp, lane,
        [&](int i, int k) { return -__ldg(gE + i + k * n); },
        [&](int k, int j) { return P[k + j * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) {
          if (j < n) gfbg[(rA + i) * n + j] = v;
          else ff[rA + i] = v;
        });
    }
    cta_sync<NT>();
    PH(11);
    // (11) [Pmat | pvec] = [Qhat | qhat] + Shat [K | kff]                                 (:249-250)
    strip_gemm<n, na, s, NW, false, false>(warp, lane,
        [&](int i, int k) { return Sh[i + k * n]; },
        [&](int k, int j) { return j < n ? fb[k * n + j] : ff[k]; },
        [&](int i, int j) { return j < n ? Qh[i + j * n] : qv[i]; },
        [&](int i, int j, double v) { P[i + j * n] = v; });
    PH(12);
    store_gains(t);
    store_value(t);
    fence_proxy_async();
    cta_sync<NT>();
    PH(13);
  }
  PH(14);

  // ---- initial stage: BK of [Pmat0 G0^T; G0 -mudyn I]        (proximal-riccati.hxx:44-62) ----
  {
    if (tid == 0) tma_store_wait_read();
    for (int e = tid; e < n * n; e += NT)
      if (!isfinite(P[e])) flags |= GARB200_FLAG_NONFINITE;
    cta_sync<NT>();
    const int c0 = p.nc0, s0 = n + c0;
    const double *G0 = p.G0 + (size_t)b * c0 * n, *g0 = p.g0 + (size_t)b * c0;
    double *off = p.kkt0_ff + (size_t)b * s0;
    // G0 == I exactly: x0 = -g0 + mudyn lbd0, (I + mudyn P0) lbd0 = P0 g0 - p0 (see sweep_fast.cu)
    int isI = (c0 == n);
    if (isI)
      for (int e = tid; e < n * n; e += NT) {
        const int i = e % n, j = e / n;
        if (G0[e] != (i == j ? 1.0 : 0.0)) isI = 0;
      }
    isI = __syncthreads_and(isI);
    if (isI) {
      double fro = 0.0;
      for (int e = tid; e < n * n; e += NT) fro = fma(P[e], P[e], fro);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) fro += __shfl_xor_sync(0xffffffffu, fro, o);
      if (lane == 0) red[warp] = fro;
      gemv4<n, n, NT>(tid, [&](int i, int j) { return P[i + j * n]; }, [&](int j) { return g0[j]; },
                      [&](int i) { return -pv[i]; }, [&](int i, double v) { x1v[i] = v; });
      cta_sync<NT>();
      fro = 0.0;
#pragma unroll
      for (int w = 0; w < NW; w++) fro += red[w];
      const double *lam = x1v;
      if (mudyn * sqrt(fro) <= 1e-9) {
        gemv4<n, n, NT>(tid, [&](int i, int j) { return -mudyn * P[i + j * n]; }, [&](int j) { return x1v[j]; },
                        [&](int i) { return x1v[i]; }, [&](int i, double v) { dinv[i] = v; });
        lam = dinv;
      } else {
        for (int e = tid; e < n * n; e += NT) {
          const int i = e % n, j = e / n;
          if (i >= j) X1[e] = (i == j ? 1.0 : 0.0) + mudyn * P[e];
        }
        cta_sync<NT>();
        if (!ldlt_solve_blocked<n, 1, NT>(tid, X1, RhsColMajor<n>{x1v}, Gs, dinv)) flags |= GARB200_FLAG_LLT_NOT_PD;
      }
      cta_sync<NT>();
      for (int i = tid; i < n; i += NT) { off[i] = fma(mudyn, lam[i], -g0[i]); off[n + i] = lam[i]; }
    } else {
      const int ld0 = s0 | 1;                        // odd leading dimension: conflict-free rows and columns
      double *K0 = sm;                               // overlays X1 .. P (see BigPlan)
      double *ff0 = Gs;
      double *wk0 = Gs + ((s0 + 1) & ~1);
      for (int i = tid; i < s0; i += NT) ff0[i] = (i < n) ? -pv[i] : -g0[i - n];
      for (int e = tid; e < s0 * n; e += NT) {       // columns < n: [P ; G0], end before the P block
        const int i = e % s0, j = e / s0;
        K0[i + j * ld0] = (i < n) ? P[i + j * n] : G0[(i - n) + j * c0];
      }
      cta_sync<NT>();
      for (int e = tid; e < s0 * c0; e += NT) {      // columns >= n (may overwrite P: no longer needed)
        const int i = e % s0, j = n + e / s0;
        K0[i + j * ld0] = (i == j) ? -mudyn : 0.0;
      }
      cta_sync<NT>();
      if (!bk_factor_cta<NT>(tid, s0, K0, ld0, kpiv, ksub, wk0, kpiv + s0)) flags |= GARB200_FLAG_BK_SINGULAR;
      if (warp == 0) bk_solve_vec_warp<(2 * n + 31) / 32>(lane, s0, K0, ld0, kpiv, ksub, ff0, reinterpret_cast<int *>(wk0));
      cta_sync<NT>();
      for (int i = tid; i < s0; i += NT) off[i] = ff0[i];
    }
  }
  if (flags) atomicOr(reinterpret_cast<int *>(p.flags) + b, flags);
  if (tid == 0) tma_store_wait_all();
  PH(15);
  if constexpr (PROF) {
    if (tid == 0 && p.phase_cycles)
#pragma unroll
      for (int i = 0; i < 16; i++) atomicAdd(reinterpret_cast<unsigned long long *>(p.phase_cycles) + i, (unsigned long long)ph_acc[i]);
  }
}

template <int NX, int NU, int NW, bool EX = false>
static int launch_big_t(const DevParams &p, cudaStream_t st) {
  using PL = BigPlan<NX, NU>;
  if (p.KL.stride != PL::kstride || p.GL.stride != PL::gstride || p.KL.E != PL::oE || p.KL.f != PL::of ||
      p.KL.A != PL::oA || p.KL.R != PL::oR || p.GL.fb != PL::gfb)
    return -3;
  const long long total = (long long)p.batch * p.T;
  int launched = 0;
  const bool lu_prepass = p.tune.einv_mode == 2;
  if (EX) {
    // explicit dynamics: no E^{-1}, no pre-pass
  } else if (total > 0 && p.tune.einv_mode != 2 && p.tune.einv_mode != 4) {   // default: tiles in registers, mirror in shared memory
    // (einv_mode 4: the matrix-in-shared-memory version below, 14.9 vs 11.1 ms on c5; 2: LU in shared memory; 10 warps
    //  per matrix measured slower: 57.7 vs 54.8 ms backward)
    constexpr int EW = 5;
    constexpr size_t esm = (size_t)GjCtaRegs<NX, EW>::SMEM * sizeof(double) + (size_t)GjCtaRegs<NX, EW>::SINTS * sizeof(int);
    if (cudaFuncSetAttribute(einv_cta_regs_kernel<NX, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)esm) != cudaSuccess) return -2;
    einv_cta_regs_kernel<NX, EW><<<(unsigned)total, 32 * EW, esm, st>>>(p);
    if (cudaGetLastError() != cudaSuccess) return -2;
    launched++;
  } else if (total > 0 && !lu_prepass) {
    constexpr int EW = 5;
    constexpr size_t esm = (size_t)GjCta<NX>::SMEM * sizeof(double) + (size_t)GjCta<NX>::SINTS * sizeof(int);
    if (cudaFuncSetAttribute(einv_cta_dmma_kernel<NX, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)esm) != cudaSuccess) return -2;
    einv_cta_dmma_kernel<NX, EW><<<(unsigned)total, 32 * EW, esm, st>>>(p);
    if (cudaGetLastError() != cudaSuccess) return -2;
    launched++;
  } else if (total > 0) {
    constexpr int ENT = 256;
    const size_t esm = (size_t)3 * NX * NX * sizeof(double) + (size_t)3 * NX * sizeof(int) + 16;
    if (cudaFuncSetAttribute(einv_cta_kernel<ENT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)esm) != cudaSuccess) return -2;
    einv_cta_kernel<ENT><<<(unsigned)total, ENT, esm, st>>>(p);
    if (cudaGetLastError() != cudaSuccess) return -2;
    launched++;
  }
  const size_t smem = (size_t)PL::total * sizeof(double);
  auto ks = p.phase_cycles ? sweep_big_kernel<NX, NU, NW, 1, true, EX> : sweep_big_kernel<NX, NU, NW, 1, false, EX>;
  if (cudaFuncSetAttribute(ks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -2;
  ks<<<p.batch, 32 * NW, smem, st>>>(p);
  if (cudaGetLastError() != cudaSuccess) return -2;
  return launched + 1;
}

bool big_path_available(int nx, int nu, int nc, int nth) { return nx == 76 && nu == 32 && nc == 0 && nth == 0; }

int launch_backward_big(const DevParams &p, cudaStream_t st) {
  if (p.nx == 76 && p.nu == 32 && p.nc == 0 && p.nth == 0)
    return p.explicit_dyn ? launch_big_t<76, 32, 10, true>(p, st) : launch_big_t<76, 32, 10>(p, st);
  return -3;
}

} // namespace garb200
