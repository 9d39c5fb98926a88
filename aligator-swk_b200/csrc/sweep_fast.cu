// sweep_fast.cu — dimension-specialised proximal Riccati backward sweep (sm_100a).
//
// Same algorithm as backward.cu (the generic kernel), restated for compile-time
// (NX, NU, NC) with nth = 0, terminal knot nu = nc = 0 — the shapes of the BASELINE configs:
//   einv_kernel    Einv_t = PartialPivLU(E_t).solve(I) for EVERY (problem, knot) in one
//                  embarrassingly parallel pre-pass (riccati-kernel.hxx:172-174).  E_t is
//                  input data, so its inversion does not belong on the sequential chain.
//   sweep_kernel   terminalSolve / stageKernelSolve / initial stage
//                  (riccati-kernel.hxx:90-150, 167-250; proximal-riccati.hxx:44-62), one CTA
//                  (NW warps) per problem, knot record + Einv staged by 1-D TMA bulk copies,
//                  all n^3 contractions on the FP64 tensor pipe (DMMA), GEMVs riding along
//                  as an extra column of the GEMM they belong to.
#include <type_traits>
#include "fast.cuh"
#include "kernels.cuh"

namespace garb200 {

// ---- pre-pass: batched inverse of E ------------------------------------------------------
// One matrix per LPM lanes (two 12x12 matrices or one 36x36 matrix per warp), the whole matrix
// in registers; E is read once from HBM and Einv written once (2 n^2 doubles of traffic).
template <int N, int WPB, int MINB = 1>
__global__ void __launch_bounds__(32 * WPB, MINB) einv_kernel(const DevParams p) {
  using GS = GjShape<N>;
  constexpr int MPW = GS::MPW, LPM = GS::LPM, NP = (N + 1) & ~1;
  __shared__ __align__(16) double s_row[WPB * MPW][NP];
  __shared__ int s_piv[WPB * MPW][NP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mi = lane / LPM, sub = lane % LPM;
  const unsigned mask = (LPM == 32) ? 0xffffffffu : (((1u << LPM) - 1u) << (LPM * mi));
  const long long total = (long long)p.batch * p.T;
  long long w = ((long long)blockIdx.x * WPB + warp) * MPW + mi;
  const bool valid = w < total;
  if (!valid) w = total - 1;      // keep the lane group busy on a duplicate, store nothing
  const int b = (int)(w / p.T), t = (int)(w % p.T);
  const double *Eg = p.knots + ((size_t)b * (p.T + 1) + phys_t(p, t)) * p.KL.stride + p.KL.E;
  double *Xg = p.einv + ((size_t)b * p.T + phys_t(p, t)) * (size_t)(N * N);
  const bool ok = gj_inverse_regs<N>(sub, mask, Eg, Xg, valid, s_row[warp * MPW + mi], s_piv[warp * MPW + mi]);
  if (!ok && sub == 0 && valid) atomicOr(reinterpret_cast<int *>(p.flags) + b, GARB200_FLAG_LU_SINGULAR);
}

// Same pre-pass on the tensor pipe (block Gauss-Jordan, fast.cuh): one warp per matrix.
template <int N, int WPB>
__global__ void __launch_bounds__(32 * WPB, 12 / WPB) einv_dmma_kernel(const DevParams p) {
  using G = GjDmma<N>;
  extern __shared__ __align__(16) double s_dyn[];
  double *s_w = s_dyn + (threadIdx.x >> 5) * G::SMEM;
  int *s_i = reinterpret_cast<int *>(s_dyn + WPB * G::SMEM) + (threadIdx.x >> 5) * G::SINTS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total = (long long)p.batch * p.T;
  long long w = (long long)blockIdx.x * WPB + warp;
  const bool valid = w < total;
  if (!valid) w = total - 1;
  const int b = (int)(w / p.T), t = (int)(w % p.T);
  const double *Eg = p.knots + ((size_t)b * (p.T + 1) + phys_t(p, t)) * p.KL.stride + p.KL.E;
  double *Xg = p.einv + ((size_t)b * p.T + phys_t(p, t)) * (size_t)(N * N);
  const bool ok = gj_inverse_dmma<N>(lane, Eg, Xg, valid, s_w, s_i);
  if (!ok && lane == 0 && valid) atomicOr(reinterpret_cast<int *>(p.flags) + b, GARB200_FLAG_LU_SINGULAR);
}

// ---- shared-memory plan of the sweep (doubles) ---------------------------------------------
// HREC: parameter dimension of the knot RECORD (leg mode: the record has no theta blocks, the
// leg-end knot supplies Gx = A^T, Gu = B^T, Gth = -mu I, gamma = f as views; parallel-solver.hpp:43-51)
// EG: E^{-1} is NOT staged in shared memory (it enters its three products as the strip-side operand
// straight from L2, see strip_gemm); the shared-memory copy of the record then has no E block.
template <int NX, int NU, int NC, int NTH = 0, int HREC = NTH, bool EG = false> struct FastPlan {
  static constexpr int n = NX, m = NU, c = NC, s = NU + NC, h = NTH, hr = HREC;
  static constexpr int esh = EG ? NX * NX : 0;      // shift of the record blocks behind E in shared memory
  static constexpr int NA = 1 + h;                      // augmented columns: the vector + the theta block
  static constexpr int na = n + NA;
  static constexpr int sp = (s + 3) & ~3;               // KKT block padded to a multiple of 4 (blocked LDL^T)
  static constexpr int r2(int x) { return (x + 1) & ~1; }
  // packed knot record (same offsets as garb200_make_knot_layout(NX, NU, NC, NTH))
  static constexpr int oQ = 0, oS = r2(oQ + n * n), oR = r2(oS + n * m), oq = r2(oR + m * m),
                       orr = r2(oq + n), oA = r2(orr + m), oB = r2(oA + n * n), oE = r2(oB + n * m),
                       of = r2(oE + n * n), oC = r2(of + n), oD = r2(oC + c * n), od = r2(oD + c * m),
                       oGth = r2(od + c), oGx = r2(oGth + hr * hr), oGu = r2(oGx + n * hr),
                       ogam = r2(oGu + m * hr), kstride = r2(ogam + hr);
  static_assert(oS == oQ + n * n && oB == oA + n * n, "[Q|S] and [A|B] must be contiguous (even n*n)");
  static constexpr int nrows = m + c + 2 * n;
  static constexpr int gff = 0, gfb = r2(nrows), gfth = r2(gfb + nrows * n), gstride = r2(gfth + nrows * h);
  // only ff, the K / Z rows of fb (operands of later steps) and fth are staged in shared memory;
  // the L and A rows of fb go from the accumulators straight to the gains record in HBM
  static constexpr int gpart = r2(gfb + s * n);
  static constexpr int gthsz = r2(nrows * h);
  static_assert(n % 2 == 0, "row-major fb rows are written as 16-byte pairs");
  // smem
  static constexpr int rec = 0;
  static constexpr int X1 = r2(rec + kstride - esh);    // n x na
  static constexpr int P = r2(X1 + n * na);             // n x na: [P | pvec | Vxt] / [Vxx | vx | Lth0] / [A_pre | tv | Yth_pre]
  static constexpr int BtV = r2(P + n * na);            // m x na
  static constexpr int kkt = r2(BtV + m * na);          // sp x sp
  static constexpr int ksub = r2(kkt + sp * sp);        // max(s, 2n) + 2
  static constexpr int KS = (sp > 2 * n ? sp : 2 * n) + 2;
  static constexpr int dinv = EG ? ksub : r2(ksub + KS); // n   (EG: ksub is aliased, see sweep_kernel)
  static constexpr int rhat = r2(dinv + n);             // 16 doubles of reduction scratch
  static constexpr int atvx = r2(rhat + 16);            // n + m
  static constexpr int G = r2(atvx + n + m);            // gains staging: [ff | fb rows K,Z] then fth
  static constexpr int Gth_s = r2(G + gpart);           // fth staging (nrows x h)
  static constexpr int Pxt = r2(Gth_s + gthsz);         // Pxttilde = -Einv^T Vxt (n x h)
  static constexpr int Vxn = r2(Pxt + n * h);           // Vxt^T Afb (h x n, transposed: n x h col-major)
  static constexpr int Vtt = r2(Vxn + n * h);           // h x h | vt h | parts h*h + h
  static constexpr int ints = r2(Vtt + 2 * (h * h + h)); // pivots: KS ints (EG: aliased, see sweep_kernel)
  static constexpr int bar = EG ? ints : r2(ints + (KS + 1) / 2);
  static constexpr int total = bar + 2;
};

// LEG: one CTA per leg of the parallel-in-time solver (parallel-solver.hxx:111-143): knots [beg, end),
// NTH = NX for the legs that end on a co-state cut (the leg-end knot is a terminal solve with nu > 0
// and the coupling blocks as its parameter block), NTH = 0 for the last leg; the value function of
// the leg's first knot goes to value slot `leg`, there is no initial stage.
// GRP > 1: GRP problems per CTA in LOCKSTEP.  Threads [g NT, (g+1) NT) solve problem blockIdx.x GRP + g
// in their own slice of shared memory with their own named barrier; once per knot all groups meet at
// barrier 0 right before the reduced-KKT solve, so that the latency-bound dependent FP64 chains of the
// factorizations run while NO group feeds the tensor pipe (a dependent DFMA queued behind the other
// CTAs' DMMAs waits ~100 cycles; alone it takes ~10), and the GEMM phases of all groups saturate the pipe
// together.  Independent CTAs (GRP = 1, 3 per SM) drift apart and pay that queueing in every chain.
// EX: explicit dynamics, E_t = -I at every knot (explicit-dynamics.hxx:32, solver-proxddp.hxx:826 — what
// ProxDDP feeds).  Then E^{-1} = -I exactly: no pre-pass, the E block is never read, product (1) is a sign
// flip, (2) the lower-symmetrisation of Pmat, (10) a copy — the same numbers as the general path, bit for bit.
template <int NX, int NU, int NC, int NTH, int NW, bool PROF, bool LEG = false, bool EG = false, int GRP = 1, bool EX = false>
__global__ void __launch_bounds__(32 * NW * GRP, GRP > 1 ? 1 : (NW >= 8 ? 2 : (NW == 5 && NTH <= 4) ? 3 : NW == 1 ? 24 : 1))
sweep_kernel(const DevParams p) {
  extern __shared__ __align__(16) double sm_cta[];
  using PL = FastPlan<NX, NU, NC, NTH, LEG ? 0 : NTH, EG>;
  static_assert(GRP == 1 || !LEG, "lockstep groups are not combined with leg mode");
  static_assert(PL::total % 2 == 0, "group slices must stay 16-byte aligned");
  const int grp = GRP > 1 ? (int)threadIdx.x / (32 * NW) : 0;
  double *sm = sm_cta + (size_t)grp * PL::total;
  static_assert(!LEG || NC == 0, "leg mode: nc = 0");
  static_assert(!EG || !LEG, "EG is not combined with leg mode");
  constexpr int n = NX, m = NU, c = NC, s = NU + NC, h = NTH, NT = 32 * NW;
  constexpr int NA = PL::NA, na = PL::na, sp = PL::sp;
  constexpr int hh = h > 0 ? h : 1;        // divisor-safe
  constexpr int rK = 0, rZ = m, rL = m + c, rA = m + c + n;
  const int tid = (int)threadIdx.x - grp * NT, warp = tid >> 5, lane = tid & 31;
  const int T = p.T;
  const int nlegs = LEG ? p.num_legs : 1;
  const int leg = LEG ? (NTH > 0 ? (int)(blockIdx.x % (unsigned)(nlegs - 1)) : nlegs - 1) : 0;
  // (a group past the end of the batch re-solves the last problem: same bytes to the same addresses)
  const int b = (LEG && NTH > 0) ? (int)(blockIdx.x / (unsigned)(nlegs - 1))
                                 : (GRP > 1 ? min((int)blockIdx.x * GRP + grp, p.batch - 1) : (int)blockIdx.x);
  int beg = 0, end = T + 1;
  if constexpr (LEG) get_work(T, leg, nlegs, beg, end);
  const int tl = end - 1;                  // last knot of this sweep (T unless a coupled leg)
  const double mudyn = p.mudyn, mueq = p.mueq;
  double *rec = sm + PL::rec;
  double *Q = rec + PL::oQ, *S = rec + PL::oS, *R = rec + PL::oR, *qv = rec + PL::oq, *rv = rec + PL::orr;
  double *A = rec + PL::oA, *B = rec + PL::oB, *Einv = rec + PL::oE, *f = rec + PL::of - PL::esh;
  double *C = rec + PL::oC - PL::esh, *D = rec + PL::oD - PL::esh, *d = rec + PL::od - PL::esh;
  double *Gth = rec + PL::oGth - PL::esh, *Gx = rec + PL::oGx - PL::esh, *Gu = rec + PL::oGu - PL::esh, *gam = rec + PL::ogam - PL::esh;
  const double *gE = nullptr;                // EG: E^{-1} of the current knot in global memory
  double *fth = sm + PL::Gth_s;          // fth staging, nrows x h row-major
  double *Pxt = sm + PL::Pxt, *Vxn = sm + PL::Vxn;
  double *Vtt = sm + PL::Vtt, *vt = Vtt + h * h, *Vttp = vt + h, *vtp = Vttp + h * h;
  double *X1 = sm + PL::X1, *P = sm + PL::P, *BtV = sm + PL::BtV, *kkt = sm + PL::kkt;
  // EG: the Bunch-Kaufman scratch (used only by the fallback paths) aliases dead blocks: P + 128.. in
  // the stage fallback (Vxx is consumed by then, the LDL^T attempt used P[0, 5 sp) itself), BtV in the
  // initial stage
  double *ksub = EG ? sm + PL::P + 128 : sm + PL::ksub;
  double *dinv = sm + PL::dinv, *atvx = sm + PL::atvx, *Gs = sm + PL::G;
  double *pv = P + n * n;         // column n of the augmented P block
  double *x1v = X1 + n * n;       // column n of the augmented X1 block
  double *btvx = BtV + m * n;     // column n of BtV: B^T vx
  double *Pth = P + n * (n + 1);  // theta columns of the P block: Vxt / Lth0 / Yth_pre
  double *BtVth = BtV + m * (n + 1);
  double *ff = Gs + PL::gff, *fb = Gs + PL::gfb;
  int *kpiv = reinterpret_cast<int *>(EG ? sm + PL::P + 128 + PL::KS : sm + PL::ints);
  static_assert(!EG || 128 + PL::KS + (PL::KS + 1) / 2 <= NX * NX, "aliased BK scratch must stay inside the dead Vxx block");
  double *red = sm + PL::rhat;    // NW doubles: cross-warp reduction scratch
  uint64_t *bar = reinterpret_cast<uint64_t *>(sm + PL::bar);
  int flags = 0;
  // optional phase timers (GARB200_PHASE_TIMERS=1): cycles of thread 0 between phase marks
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

  // two transaction barriers per knot: `bar` for what the first products need (E^{-1}; everything for the
  // terminal knot and for E = -I), `bar2` for the rest of the record ([Q..B], [f..]), first read in step (3):
  // its 30 KB arrive from L2 while the two E^{-1} products run
  uint64_t *bar2 = bar + 1;
  if (tid == 0) { mbar_init(bar, 1); mbar_init(bar2, 1); fence_mbar_init(); }
  for (int e = tid; e < PL::gpart; e += NT) Gs[e] = 0.0;
  for (int e = tid; e < PL::gthsz; e += NT) fth[e] = 0.0;
  cta_sync<NT, GRP>();
  uint32_t phase = 0, phase2 = 0;
  const size_t pbase = (size_t)b * (T + 1);
  auto issue_load = [&](int t) {      // thread 0 only
    const int pt = phys_t(p, t);
    const double *src = p.knots + (pbase + pt) * PL::kstride;
    fence_proxy_async();
    if constexpr (EX) {                 // [Q..B] and [f..]: E = -I is never read
      mbar_expect_tx(bar, (uint32_t)(PL::kstride - n * n) * 8u);
      tma_load_1d(rec, src, (uint32_t)PL::oE * 8u, bar);
      tma_load_1d(rec + (EG ? PL::oE : PL::of), src + PL::of, (uint32_t)(PL::kstride - PL::of) * 8u, bar);
      return;
    }
    if constexpr (EG) {                 // [Q..B] and [f..], the E block is never staged
      mbar_expect_tx(bar2, (uint32_t)(PL::kstride - PL::esh) * 8u);
      tma_load_1d(rec, src, (uint32_t)PL::oE * 8u, bar2);
      tma_load_1d(rec + PL::oE, src + PL::of, (uint32_t)(PL::kstride - PL::of) * 8u, bar2);
      return;
    }
    if (t == tl) {                      // terminal / leg-end knot: E is not used
      mbar_expect_tx(bar, (uint32_t)PL::kstride * 8u);
      tma_load_1d(rec, src, (uint32_t)PL::kstride * 8u, bar);
    } else {
      mbar_expect_tx(bar, (uint32_t)(n * n) * 8u);
      tma_load_1d(Einv, p.einv + ((size_t)b * T + pt) * (size_t)(n * n), (uint32_t)(n * n) * 8u, bar);
      mbar_expect_tx(bar2, (uint32_t)(PL::kstride - n * n) * 8u);
      tma_load_1d(rec, src, (uint32_t)PL::oE * 8u, bar2);
      tma_load_1d(rec + PL::of, src + PL::of, (uint32_t)(PL::kstride - PL::of) * 8u, bar2);
    }
  };
  auto store_gains = [&](int t) {     // all threads; Gs complete
    fence_proxy_async();
    cta_sync<NT, GRP>();
    if (tid == 0) {
      double *gdst = p.gains + (pbase + phys_t(p, t)) * p.GL.stride;
      tma_store_1d(gdst, Gs, (uint32_t)PL::gpart * 8u);
      if constexpr (h > 0) tma_store_1d(gdst + PL::gfth, fth, (uint32_t)PL::gthsz * 8u);
      tma_store_commit();
    }
  };
  auto store_value = [&](int t) {     // [P | pv] is the value function of knot t
    const int slot = p.store_all_values ? t : (LEG ? (t == beg ? leg : -1) : (t == 0 ? 0 : -1));
    if (slot < 0) return;
    double *v = p.values + ((size_t)b * p.nslots + slot) * p.VL.stride;
    for (int e = tid; e < n * n; e += NT) v[p.VL.Pmat + e] = P[e];
    for (int i = tid; i < n; i += NT) v[p.VL.pvec + i] = pv[i];
    if constexpr (h > 0) {
      for (int e = tid; e < n * h; e += NT) v[p.VL.Vxt + e] = Pth[e];
      for (int e = tid; e < h * h; e += NT) v[p.VL.Vtt + e] = Vtt[e];
      for (int i = tid; i < h; i += NT) v[p.VL.vt + i] = vt[i];
    }
    if (!LEG && t == 0 && p.store_all_values) {               // proximal-riccati.hxx:39-42
      for (int e = tid; e < n * n; e += NT) v[p.VL.Vxx + e] = P[e];
      for (int i = tid; i < n; i += NT) v[p.VL.vx + i] = pv[i];
    }
  };
  // leg mode, last leg: the gains records carry an (unused) fth block of the coupled legs' width
  auto zero_fth_tail = [&](int t) {
    if constexpr (LEG && h == 0) {
      double *g = p.gains + (pbase + phys_t(p, t)) * p.GL.stride;
      for (int e = PL::gfth + tid; e < p.GL.stride; e += NT) g[e] = 0.0;
    }
  };

  // ---- terminal knot (nu = nc = 0): Pmat = Q, pvec = q, empty gains  (riccati-kernel.hxx:90-150)
  if (tid == 0) issue_load(tl);
  if constexpr (EG && !EX) { mbar_wait(bar2, phase2); phase2 ^= 1; }     // (EG stages every knot through bar2)
  else { mbar_wait(bar, phase); phase ^= 1; }
  if constexpr (LEG && h > 0) {
    // ---- leg-end knot of a coupled leg: terminalSolve with nu > 0 and the parameter block
    //      Gx = A^T, Gu = B^T, Gth = -mudyn I, gamma = f     (riccati-kernel.hxx:110-149, parallel-solver.hpp:43-51)
    auto build_term = [&]() {
      for (int e = tid; e < m * m; e += NT) { const int i = e % m, j = e / m; kkt[i + j * sp] = R[i + j * m]; }
      if constexpr (sp > s) {
        constexpr int pr = sp - s > 0 ? sp - s : 1;
        for (int e = tid; e < (sp - s) * sp; e += NT) { const int i = s + e % pr, j = e / pr; kkt[i + j * sp] = (i == j) ? -1.0 : 0.0; }
      }
      for (int i = tid; i < m; i += NT) ff[rK + i] = -rv[i];
      for (int e = tid; e < m * n; e += NT) { const int i = e / n, j = e % n; fb[(rK + i) * n + j] = -S[j + i * n]; }
      for (int e = tid; e < m * h; e += NT) { const int i = e / hh, j = e % hh; fth[(rK + i) * h + j] = -B[j + i * n]; }   // -Gu = -B^T
    };
    build_term();
    cta_sync<NT, GRP>();
    if (!ldlt_solve_blocked<sp, na, NT, m, GRP>(tid, kkt, RhsGainsTh<n, h, s>{ff, fb, fth}, P, P + 4 * sp)) {
      build_term();
      cta_sync<NT, GRP>();
      if (warp == 0) {
        if (!bk_factor_warp(lane, s, kkt, sp, kpiv, ksub)) flags |= GARB200_FLAG_BK_SINGULAR;
      }
      cta_sync<NT, GRP>();
      for (int col = tid; col < na; col += NT) {
        if (col == 0) bk_solve_column(s, kkt, sp, kpiv, ksub, ff, 1);
        else if (col <= n) bk_solve_column(s, kkt, sp, kpiv, ksub, fb + (col - 1), n);
        else bk_solve_column(s, kkt, sp, kpiv, ksub, fth + (col - n - 1), h);
      }
      cta_sync<NT, GRP>();
    }
    // [Pmat | pvec] = [Q | q] + S [K | kff]                                              (:134-140)
    mma_gemm<n, n + 1, s, NW>(warp, lane,
        [&](int i, int k) { return S[i + k * n]; },
        [&](int k, int j) { return j < n ? fb[k * n + j] : ff[k]; },
        [&](int i, int j) { return j < n ? Q[i + j * n] : qv[i]; },
        [&](int i, int j, double v) { P[i + j * n] = v; });
    // Vxt = A^T + K^T B^T ; Vtt = -mudyn I + B Kth ; vt = f + B kff                       (:142-149)
    for (int e = tid; e < n * h; e += NT) {
      const int i = e % n, j = e / n;
      double v = A[j + i * n];
#pragma unroll
      for (int k = 0; k < m; k++) v = fma(fb[(rK + k) * n + i], B[j + k * n], v);
      Pth[e] = v;
    }
    for (int e = tid; e < h * h + h; e += NT) {
      if (e < h * h) {
        const int a = e % hh, bcol = e / hh;
        double v = (a == bcol) ? -mudyn : 0.0;
#pragma unroll
        for (int k = 0; k < m; k++) v = fma(B[a + k * n], fth[(rK + k) * h + bcol], v);
        Vtt[e] = v;
      } else {
        const int a = e - h * h;
        double v = f[a];
#pragma unroll
        for (int k = 0; k < m; k++) v = fma(B[a + k * n], ff[rK + k], v);
        vt[a] = v;
      }
    }
  } else {
    for (int e = tid; e < n * n; e += NT) P[e] = Q[e];
    for (int i = tid; i < n; i += NT) pv[i] = qv[i];
    if constexpr (h > 0) {                                  // Vxt = Gx, Vtt = Gth, vt = gamma  (:142-149, nu = 0)
      for (int e = tid; e < n * h; e += NT) Pth[e] = Gx[e];
      for (int e = tid; e < h * h; e += NT) Vtt[e] = Gth[e];
      for (int i = tid; i < h; i += NT) vt[i] = gam[i];
    }
  }
  store_gains(tl);
  {
    double *gT = p.gains + (pbase + phys_t(p, tl)) * p.GL.stride;
    for (int e = PL::gpart + tid; e < PL::gfth; e += NT) gT[e] = 0.0;
  }
  zero_fth_tail(tl);
  store_value(tl);
  fence_proxy_async();
  cta_sync<NT, GRP>();

#pragma unroll 1
  for (int t = tl - 1; t >= beg; t--) {
    if (tid == 0) {
      issue_load(t);
      if (t > beg) {          // next knot's record and E^{-1} towards L2 while this one is solved
        const int pn = phys_t(p, t - 1);
        l2_prefetch(p.knots + (pbase + pn) * PL::kstride, (uint32_t)PL::kstride * 8u);
        if constexpr (!EX) l2_prefetch(p.einv + ((size_t)b * T + pn) * (size_t)(n * n), (uint32_t)(n * n) * 8u);
      }
      tma_store_wait_read();
    }
    if constexpr (EG && !EX) gE = p.einv + ((size_t)b * T + phys_t(p, t)) * (size_t)(n * n);
    if constexpr (EX || !EG) { mbar_wait(bar, phase); phase ^= 1; }
    cta_sync<NT, GRP>();
    PH(0);
    double fro = 0.0;
    if constexpr (EX) {
      // E^{-1} = -I: EinvP = -Pmat and -ptilde = -pvec are never formed; Ptilde = (Pmat) lower-symmetrised
      // (:179-180, exact), Pxttilde = -Einv^T Vxt = Vxt                                      (:261-262)
      for (int e = tid; e < n * n; e += NT) {
        const int i = e % n, j = e / n;
        if (i >= j) {
          const double v = P[e];
          if (i > j) P[j + i * n] = v;
          fro = fma(i == j ? v : 2.0 * v, v, fro);
        }
      }
      if constexpr (h > 0)
        for (int e = tid; e < n * h; e += NT) Pxt[e] = Pth[e];
      PH(1);
    } else {
      // (1) [X1 | x1v] = Einv^T [P | pvec]          EinvP, -ptilde      (:175-178)
      //     theta columns: Pxttilde = -Einv^T Vxt -> Pxt                                   (:261-262)
      if constexpr (EG)
        strip_gemm<n, na, n, NW, true, false>(warp, lane,
            [&](int i, int k) { return __ldg(gE + k + i * n); },
            [&](int k, int j) { return P[k + j * n]; },
            [](int, int) { return 0.0; },
            [&](int i, int j, double v) { if (j <= n) X1[i + j * n] = v; else Pxt[i + (j - n - 1) * n] = -v; });
      else
      mma_gemm<n, na, n, NW>(warp, lane,
          [&](int i, int k) { return Einv[k + i * n]; },
          [&](int k, int j) { return P[k + j * n]; },
          [](int, int) { return 0.0; },
          [&](int i, int j, double v) { if (j <= n) X1[i + j * n] = v; else Pxt[i + (j - n - 1) * n] = -v; });
      cta_sync<NT, GRP>();
      PH(1);
      // (2) Ptilde = X1 * Einv (lower triangle kept, mirrored); ||Ptilde||_F^2 on the fly  (:179-180)
      if constexpr (EG)      // transposed: Ptilde^T = Einv^T X1^T, whose upper triangle is the reference's lower one
        strip_gemm<n, n, n, NW, true, false>(warp, lane,
            [&](int i, int k) { return __ldg(gE + k + i * n); },
            [&](int k, int j) { return X1[j + k * n]; },
            [](int, int) { return 0.0; },
            [&](int i, int j, double v) {
              if (i <= j) { P[j + i * n] = v; P[i + j * n] = v; fro = fma(i == j ? v : 2.0 * v, v, fro); }
            });
      else
      mma_gemm_lower<n, n, NW>(warp, lane,
          [&](int i, int k) { return X1[i + k * n]; },
          [&](int k, int j) { return Einv[k + j * n]; },
          [](int, int) { return 0.0; },
          [&](int i, int j, double v) {
            if (i >= j) { P[i + j * n] = v; P[j + i * n] = v; fro = fma(i == j ? v : 2.0 * v, v, fro); }
          });
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) fro += __shfl_xor_sync(0xffffffffu, fro, o);
    if (lane == 0) red[warp] = fro;
    cta_sync<NT, GRP>();
    PH(2);
    fro = 0.0;
#pragma unroll
    for (int w = 0; w < NW; w++) fro += red[w];
    // Lambda = I + mudyn Ptilde.  When mudyn ||Ptilde||_F is small the Neumann series
    //   Lambda^{-1} X = X - mudyn Ptilde X + (mudyn Ptilde)^2 X - ...     evaluated as  Y <- X - mudyn Ptilde Y
    // is one to three products on the tensor pipe; otherwise the blocked LDL^T solve runs.  Uniform over
    // the CTA (same `fro` in every thread).
    // Order of the series: x = mudyn ||Ptilde||_F bounds the spectral radius of mudyn Ptilde, the order-k
    // partial sum has relative truncation error <= x^(k+1) <= 1e-18 (below the rounding of the LLT solve it
    // replaces): k = 1 for x <= 1e-9 (implicit dynamics with |E| >> 1), k = 2 for x <= 1e-6, k = 3 for
    // x <= 3e-5 (explicit dynamics, E = -I: Ptilde = Pmat); beyond that the LDL^T solve runs.
    const double xnorm = mudyn * sqrt(fro);
    const int order = xnorm <= 1e-9 ? 1 : xnorm <= 1e-6 ? 2 : xnorm <= 3e-5 ? 3 : 0;
    const bool neumann = order > 0;
    if constexpr (!EX) { mbar_wait(bar2, phase2); phase2 ^= 1; }     // the rest of the record is here
    // (3) w = ptilde + Ptilde f -> pv                                                   (:182-187)
    gemv4<n, n, NT>(tid, [&](int i, int j) { return P[i + j * n]; }, [&](int j) { return f[j]; },
                    [&](int i) { return EX ? pv[i] : -x1v[i]; }, [&](int i, double v) { pv[i] = v; });   // ptilde = -x1v (EX: = pvec)
    if constexpr (h > 0 && !EX)
      for (int e = tid; e < n * h; e += NT) Pth[e] = Pxt[e];     // right-hand sides of Lth0 = Lambda^{-1} Pxttilde
    if (!neumann) {
      for (int e = tid; e < n * n; e += NT) {
        const int i = e % n, j = e / n;
        if (i >= j) X1[e] = (i == j ? 1.0 : 0.0) + mudyn * P[e];
      }
    }
    cta_sync<NT, GRP>();
    PH(3);
    // (4) [Vxx | vx] = Lambda^{-1} [Ptilde | w]                                      (:185-190)
    //     Neumann form: the product lands in the X1 block and STAYS there (no copy back): from here to
    //     step (10) `VX` is the block holding [Vxx | vx | Lth0] (later [A_pre | tv | Yth_pre]) and `AT`
    //     the other one, which receives AtV.  LDL^T form: solved in place in the P block.
    if (neumann) {
      mma_gemm<n, na, n, NW>(warp, lane,      // (lower-triangle tiles only was measured slower: it loses the strip-mode fragment reuse)
          [&](int i, int k) { return -mudyn * P[i + k * n]; },
          [&](int k, int j) { return P[k + j * n]; },
          [&](int i, int j) { return P[i + j * n]; },
          [&](int i, int j, double v) {                       // last product: lower triangle kept and mirrored in the epilogue
            if (order > 1 || j >= n || i >= j) X1[i + j * n] = v;
            if (order == 1 && j < n && i > j) X1[j + i * n] = v;
          });
#pragma unroll 1
      for (int it = 1; it < order; it++) {                    // Y <- [Ptilde | w] - mudyn Ptilde Y, in place (column strips)
        cta_sync<NT, GRP>();
        mma_gemm_colstrips<n, na, n, NW>(warp, lane,
            [&](int i, int k) { return -mudyn * P[i + k * n]; },
            [&](int k, int j) { return X1[k + j * n]; },
            [&](int i, int j) { return P[i + j * n]; },
            [&](int i, int j, double v) { X1[i + j * n] = v; });
      }
      if (order > 1) {                                        // selfadjointView<Lower> of the result (:190)
        cta_sync<NT, GRP>();
        mirror_lower_ct<n, NT>(tid, X1);
      }
      PH(4);
    } else {
      if (!ldlt_solve_blocked<n, na, NT, n, GRP>(tid, X1, RhsColMajor<n>{P}, Gs, dinv)) flags |= GARB200_FLAG_LLT_NOT_PD;
      PH(4);
      mirror_lower_ct<n, NT>(tid, P);
    }
    double *const VX = neumann ? X1 : P, *const AT = neumann ? P : X1;
    double *const vxv = VX + n * n, *const VXth = VX + n * (n + 1);
    cta_sync<NT, GRP>();
    PH(5);
    if (p.store_all_values) {
      double *v = p.values + ((size_t)b * p.nslots + (t + 1)) * p.VL.stride;
      for (int e = tid; e < n * n; e += NT) v[p.VL.Vxx + e] = VX[e];
      for (int i = tid; i < n; i += NT) v[p.VL.vx + i] = vxv[i];
    }
    // (5) [AtV ; BtV] = [A | B]^T [Vxx | vx]: rows < n -> AT (+ A^T vx), rows >= n -> BtV (+ B^T vx)
    mma_gemm<n + m, na, n, NW>(warp, lane,
        [&](int i, int k) { return A[k + i * n]; },           // [A | B] is one n x (n+m) col-major block
        [&](int k, int j) { return VX[k + j * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) {
          if (i < n) { if (j < n) AT[i + j * n] = v; else if (j == n) atvx[i] = v; }   // A^T Lth0 is not needed
          else BtV[(i - n) + j * m] = v;                                             // B^T [Vxx | vx | Lth0]
        });
    cta_sync<NT, GRP>();
    PH(6);
    // (6) [Qhat | Shat] = [Q | S] + AtV [A | B] (in place)                                  (:195-199)
    //     on the first NWG warps, with -Shat^T going straight into the K rows of the right-hand side;
    // (7a) meanwhile the LAST warp builds the reduced KKT matrix [Rhat D^T; D -mueq I] (Rhat = R + BtV B)
    //     and factorises it (blocked LDL^T in natural order, inertia + rank check) — a dependent FP64
    //     chain that needs only BtV, so it runs beside the big GEMM instead of after it.       (:202-208)
    constexpr int NWG = NW > 1 ? NW - 1 : 1;          // GEMM warps
    constexpr int FW = NW - 1;                         // factor warp (NW == 1: the one warp does both in turn)
    int *kkt_ok = reinterpret_cast<int *>(red + NW);   // red has 16 doubles, NW <= 8
    auto kkt_matrix = [&](auto nwc, int w, int t0, int nt) {   // by decltype(nwc)::value warps (this one is w) = nt threads (this one is t0)
      mma_gemm<m, m, n, decltype(nwc)::value>(w, lane,
          [&](int i, int k) { return BtV[i + k * m]; },
          [&](int k, int j) { return B[k + j * n]; },
          [&](int i, int j) { return R[i + j * m]; },
          [&](int i, int j, double v) { kkt[i + j * sp] = v; });
      if constexpr (c > 0) {
        constexpr int cc = c > 0 ? c : 1;
        for (int e = t0; e < c * s; e += nt) {
          const int i = e % cc, j = e / cc;
          kkt[(m + i) + j * sp] = (j < m) ? D[i + j * c] : ((j - m == i) ? -mueq : 0.0);
        }
      }
      if constexpr (sp > s) {        // padding rows of the blocked LDL^T: -I, decoupled
        constexpr int pr = sp - s > 0 ? sp - s : 1;
        for (int e = t0; e < (sp - s) * sp; e += nt) {
          const int i = s + e % pr, j = e / pr;
          kkt[i + j * sp] = (i == j) ? -1.0 : 0.0;
        }
      }
    };
    auto kkt_rhs = [&](int t0, int nt) {     // [-rhat | . ; -d | -C], theta columns [-Guhat ; 0]     (:210-226, 264-272)
      for (int i = t0; i < m; i += nt) ff[rK + i] = -(rv[i] + btvx[i]);          // -rhat
      if constexpr (c > 0) {
        constexpr int cc = c > 0 ? c : 1;
        for (int i = t0; i < c; i += nt) ff[rZ + i] = -d[i];
        for (int e = t0; e < c * n; e += nt) { const int i = e % cc, j = e / cc; fb[(rZ + i) * n + j] = -C[i + j * c]; }
      }
      if constexpr (h > 0) {
        for (int e = t0; e < m * h; e += nt) { const int i = e / hh, j = e % hh; fth[(rK + i) * h + j] = -((LEG ? 0.0 : Gu[i + j * m]) + BtVth[i + j * m]); }
        for (int e = t0; e < c * h; e += nt) fth[rZ * h + e] = 0.0;
      }
    };
    auto build_rhs_K = [&]() {   // needs Shat complete (fallback path only)
      for (int e = tid; e < m * n; e += NT) { const int i = e / n, j = e % n; fb[(rK + i) * n + j] = -S[j + i * n]; }
    };
    if (NW == 1 || warp < NWG) {
      mma_gemm<n, n + m, n, NWG>(warp, lane,
          [&](int i, int k) { return AT[i + k * n]; },
          [&](int k, int j) { return A[k + j * n]; },
          [&](int i, int j) { return Q[i + j * n]; },           // [Q | S] contiguous as well
          [&](int i, int j, double v) {
            Q[i + j * n] = v;
            if (j >= n) fb[(rK + (j - n)) * n + i] = -v;        // -Shat^T: right-hand side of K
          });
      kkt_rhs(tid, 32 * NWG);
      for (int i = tid; i < n; i += 32 * NWG) qv[i] += atvx[i];                   // qhat
    }
    if (warp == FW) {
      kkt_matrix(std::integral_constant<int, 1>{}, 0, lane, 32);
      __syncwarp();
      const bool okf = ldlt_factor_warp<sp, m>(lane, kkt, VX, dinv);    // VX[0, 4 sp): Vxx is consumed
      if (lane == 0) *kkt_ok = okf ? 1 : 0;
    }
    PH(7);
    if constexpr (GRP > 1) __syncthreads();
    else cta_sync<NT, GRP>();
    PH(8);
    // (7b) [K kff; Z zff] = KKT^{-1} rhs: one right-hand-side column per thread               (:210-228)
    {
      const bool okk = *kkt_ok != 0;
      if (okk) {
        // small systems (12 x 12, 4 x 4): one lane group per column; larger ones (constraints: 20, 24): blocked on
        // the tensor pipe (measured: c1b 21.95 vs 22.09 ms with lanes / DMMA, c3 68.3 vs 65.1 ms)
        if constexpr (sp >= 16) ldlt_solve_dmma<sp, na, NT>(tid, kkt, dinv, RhsGainsTh<n, h, s>{ff, fb, fth});
        else ldlt_solve_lanes<sp, na, NT>(tid, kkt, dinv, RhsGainsTh<n, h, s>{ff, fb, fth});
        cta_sync<NT, GRP>();
      } else {                        // uniform: rebuild and factorise with the reference's Bunch-Kaufman pivoting
        kkt_matrix(std::integral_constant<int, NW>{}, warp, tid, NT);
        kkt_rhs(tid, NT);
        build_rhs_K();
        cta_sync<NT, GRP>();
        double *const ksb = EG ? VX + 128 : ksub;                 // EG: scratch inside the dead Vxx block
        int *const kpv = EG ? reinterpret_cast<int *>(VX + 128 + PL::KS) : kpiv;
        if (warp == 0) {
          if (!bk_factor_warp(lane, s, kkt, sp, kpv, ksb)) flags |= GARB200_FLAG_BK_SINGULAR;
        }
        cta_sync<NT, GRP>();
        for (int col = tid; col < na; col += NT) {
          if (col == 0) bk_solve_column(s, kkt, sp, kpv, ksb, ff, 1);
          else if (col <= n) bk_solve_column(s, kkt, sp, kpv, ksb, fb + (col - 1), n);
          else bk_solve_column(s, kkt, sp, kpv, ksb, fth + (col - n - 1), h);
        }
        cta_sync<NT, GRP>();
      }
    }
    PH(9);
    // (8) lff = vx + BtV^T kff ; tv = f + B kff - mudyn lff -> pv slot                  (:231-236)
    for (int i = tid; i < n; i += NT) {
      double l = vxv[i], tb = f[i];
#pragma unroll
      for (int k = 0; k < m; k++) { l = fma(BtV[k + i * m], ff[rK + k], l); tb = fma(B[i + k * n], ff[rK + k], tb); }
      ff[rL + i] = l;
      vxv[i] = tb - mudyn * l;
    }
    if constexpr (h > 0) {          // Lth = Lth0 + BtV^T Kth ; Yth_pre = B Kth - mudyn Lth        (:274-280)
      for (int e = tid; e < n * h; e += NT) {
        const int i = e % n, j = e / n;
        double l = VXth[i + j * n], tb = 0.0;
#pragma unroll
        for (int k = 0; k < m; k++) {
          const double kth = fth[(rK + k) * h + j];
          l = fma(BtV[k + i * m], kth, l);
          tb = fma(B[i + k * n], kth, tb);
        }
        fth[(rL + i) * h + j] = l;
        VXth[i + j * n] = tb - mudyn * l;
      }
    }
    // (9) L = AtV^T + BtV^T K  (Vxx A == AtV^T: Vxx is exactly symmetric) -> fb rows rL     (:238-239)
    //     A_pre = A + B K - mudyn L -> P block                                            (:241-242)
    double *gfbg = p.gains + (pbase + phys_t(p, t)) * p.GL.stride + PL::gfb;   // fb rows of this knot in HBM
    {
      constexpr int TM = (n + 7) / 8, TN = (n + 7) / 8, KSt = (m + 3) / 4, NTILE = TM * TN;
      const int g = lane >> 2, q = lane & 3;
      if constexpr (TM % NW == 0) {        // row strips: one A-side fragment pair feeds the whole strip
        constexpr int SR = TM / NW;
#pragma unroll
        for (int r = 0; r < SR; r++) {
          const int ic = (warp + NW * r) * 8 + g, ia = min(ic, n - 1);
          double l0[TN], l1[TN], a0[TN], a1[TN];
          int jbk[TN];
#pragma unroll
          for (int u = 0; u < TN; u++) {
            jbk[u] = min(u * 8 + g, n - 1);
            const int jc = u * 8 + 2 * q, j0 = min(jc, n - 1), j1 = min(jc + 1, n - 1);
            l0[u] = AT[j0 + ia * n]; l1[u] = AT[j1 + ia * n];
            a0[u] = A[ia + j0 * n];  a1[u] = A[ia + j1 * n];
          }
#pragma unroll
          for (int ks = 0; ks < KSt; ks++) {
            const int k = ks * 4 + q;
            const bool kv = (m % 4 == 0) || (k < m);
            const int kc = min(k, m - 1);
            const double aL = kv ? BtV[kc + ia * m] : 0.0, aB = kv ? B[ia + kc * n] : 0.0;
            double bvK[TN];
#pragma unroll
            for (int u = 0; u < TN; u++) bvK[u] = kv ? fb[(rK + kc) * n + jbk[u]] : 0.0;
#pragma unroll
            for (int u = 0; u < TN; u++) { dmma_884(l0[u], l1[u], aL, bvK[u]); dmma_884(a0[u], a1[u], aB, bvK[u]); }
          }
#pragma unroll
          for (int u = 0; u < TN; u++) {
            const int jc = u * 8 + 2 * q;
            if (ic < n && jc < n) {
              *reinterpret_cast<double2 *>(gfbg + (rL + ic) * n + jc) = make_double2(l0[u], l1[u]);
              VX[ic + jc * n] = a0[u] - mudyn * l0[u];
              VX[ic + (jc + 1) * n] = a1[u] - mudyn * l1[u];
            }
          }
        }
      } else {
      constexpr int NSLOT = (NTILE + NW - 1) / NW;
      double l0[NSLOT], l1[NSLOT], a0[NSLOT], a1[NSLOT];
      int ia[NSLOT], jbk[NSLOT];
#pragma unroll
      for (int u = 0; u < NSLOT; u++) {
        const int tt = min(warp + NW * u, NTILE - 1);
        const int ti = tt % TM, tj = tt / TM;
        ia[u] = min(ti * 8 + g, n - 1);
        jbk[u] = min(tj * 8 + g, n - 1);
        const int jc = tj * 8 + 2 * q, j0 = min(jc, n - 1), j1 = min(jc + 1, n - 1);
        l0[u] = AT[j0 + ia[u] * n]; l1[u] = AT[j1 + ia[u] * n];       // AtV^T
        a0[u] = A[ia[u] + j0 * n];  a1[u] = A[ia[u] + j1 * n];
      }
#pragma unroll
      for (int ks = 0; ks < KSt; ks++) {
        const int k = ks * 4 + q;
        const bool kv = (m % 4 == 0) || (k < m);
        const int kc = min(k, m - 1);
        double bvK[NSLOT], aL[NSLOT], aB[NSLOT];
#pragma unroll
        for (int u = 0; u < NSLOT; u++) {
          bvK[u] = kv ? fb[(rK + kc) * n + jbk[u]] : 0.0;
          aL[u] = kv ? BtV[kc + ia[u] * m] : 0.0;
          aB[u] = kv ? B[ia[u] + kc * n] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < NSLOT; u++) {
          dmma_884(l0[u], l1[u], aL[u], bvK[u]);
          dmma_884(a0[u], a1[u], aB[u], bvK[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < NSLOT; u++) {
        const int tt = warp + NW * u;
        if (tt < NTILE) {
          const int ti = tt % TM, tj = tt / TM;
          const int ic = ti * 8 + g, jc = tj * 8 + 2 * q;
          if (ic < n && jc < n) {
            *reinterpret_cast<double2 *>(gfbg + (rL + ic) * n + jc) = make_double2(l0[u], l1[u]);
            VX[ic + jc * n] = a0[u] - mudyn * l0[u];
            VX[ic + (jc + 1) * n] = a1[u] - mudyn * l1[u];
          }
        }
      }
      }
    }
    cta_sync<NT, GRP>();
    PH(10);
    // (10) [Afb | yff] = -Einv [A_pre | tv]                                              (:243-245)
    //      rows n..n+h-1 of the same product: Pxttilde^T [A_pre | tv | Yth_pre] = Vxt^T [Afb | yff | Yth]
    //      (Vxt^T Einv^{-1}... = Pxttilde^T exactly), the three Vxt-contractions of :286-298
    if constexpr (EX) {
      // -E^{-1} = I: [Afb | yff | Yth] = [A_pre | tv | Yth_pre], a (transposing) copy into the gains record
      for (int e = tid; e < n * n; e += NT) { const int i = e / n, j = e % n; gfbg[(rA + i) * n + j] = VX[i + j * n]; }
      for (int i = tid; i < n; i += NT) ff[rA + i] = vxv[i];
      if constexpr (h > 0) {
        for (int e = tid; e < n * h; e += NT) { const int i = e / hh, j = e % hh; fth[(rA + i) * h + j] = VXth[i + j * n]; }
        mma_gemm<h, na, n, NW>(warp, lane,                  // the Vxt contractions: Pxttilde^T [A_pre | tv | Yth_pre]
            [&](int i, int k) { return Pxt[k + i * n]; },
            [&](int k, int j) { return VX[k + j * n]; },
            [](int, int) { return 0.0; },
            [&](int a, int j, double v) {
              if (j < n) Vxn[j + a * n] = v;
              else if (j == n) vtp[a] = v;
              else Vttp[a + (j - n - 1) * h] = v;
            });
      }
    } else
    if constexpr (EG)
      strip_gemm<n + h, na, n, NW, true, false>(warp, lane,
        [&](int i, int k) { return i < n ? -__ldg(gE + i + k * n) : Pxt[k + (i - n) * n]; },
        [&](int k, int j) { return VX[k + j * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) {
          if (i < n) {
            if (j < n) gfbg[(rA + i) * n + j] = v;
            else if (j == n) ff[rA + i] = v;
            else fth[(rA + i) * h + (j - n - 1)] = v;
          } else {
            const int a = i - n;
            if (j < n) Vxn[j + a * n] = v;
            else if (j == n) vtp[a] = v;
            else Vttp[a + (j - n - 1) * h] = v;
          }
        });
    else
    mma_gemm<n + h, na, n, NW>(warp, lane,
        [&](int i, int k) { return i < n ? -Einv[i + k * n] : Pxt[k + (i - n) * n]; },
        [&](int k, int j) { return VX[k + j * n]; },
        [](int, int) { return 0.0; },
        [&](int i, int j, double v) {
          if (i < n) {
            if (j < n) gfbg[(rA + i) * n + j] = v;
            else if (j == n) ff[rA + i] = v;
            else fth[(rA + i) * h + (j - n - 1)] = v;
          } else {
            const int a = i - n;
            if (j < n) Vxn[j + a * n] = v;
            else if (j == n) vtp[a] = v;
            else Vttp[a + (j - n - 1) * h] = v;
          }
        });
    cta_sync<NT, GRP>();
    PH(11);
    // (11) [Pmat | pvec] = [Qhat | qhat] + Shat [K | kff] + C^T [Z | zff]                  (:249-250)
    mma_gemm<n, n + 1, s, NW>(warp, lane,
        [&](int i, int k) { return k < m ? S[i + k * n] : C[(k - m) + i * c]; },
        [&](int k, int j) { return j < n ? fb[k * n + j] : ff[k]; },
        [&](int i, int j) { return j < n ? Q[i + j * n] : qv[i]; },
        [&](int i, int j, double v) { P[i + j * n] = v; });
    if constexpr (h > 0) {
      // Vxt = Gx + K^T Gu + Afb^T Vxt+ ; vt += gamma + Gu^T kff + Vxt+^T yff ;
      // Vtt += Gth + Gu^T Kth + Vxt+^T Yth                                                (:286-298)
      for (int e = tid; e < n * h; e += NT) {
        const int i = e % n, j = e / n;
        double v = Vxn[e];
        if constexpr (!LEG) {
          v += Gx[e];
#pragma unroll
          for (int k = 0; k < m; k++) v = fma(fb[(rK + k) * n + i], Gu[k + j * m], v);
        }
        Pth[e] = v;
      }
      for (int e = tid; e < h * h + h; e += NT) {
        if (e < h * h) {
          const int a = e % hh, bcol = e / hh;
          double v = Vtt[e] + Vttp[e];
          if constexpr (!LEG) {
            v += Gth[e];
#pragma unroll
            for (int k = 0; k < m; k++) v = fma(Gu[k + a * m], fth[(rK + k) * h + bcol], v);
          }
          Vtt[e] = v;
        } else {
          const int a = e - h * h;
          double v = vt[a] + vtp[a];
          if constexpr (!LEG) {
            v += gam[a];
#pragma unroll
            for (int k = 0; k < m; k++) v = fma(Gu[k + a * m], ff[rK + k], v);
          }
          vt[a] = v;
        }
      }
    }
    PH(12);
    store_gains(t);
    zero_fth_tail(t);
    store_value(t);
    fence_proxy_async();
    cta_sync<NT, GRP>();
    PH(13);
  }
  PH(14);

  // ---- initial stage: BK of [Pmat0 G0^T; G0 -mudyn I]        (proximal-riccati.hxx:44-62) ----
  if constexpr (EG) { ksub = BtV; kpiv = reinterpret_cast<int *>(BtV + PL::KS); }   // P is live here, BtV is dead
  if constexpr (!LEG) {
    if (tid == 0) tma_store_wait_read();
    cta_sync<NT, GRP>();
    const int c0 = p.nc0, s0 = n + c0;
    const double *G0 = p.G0 + (size_t)b * c0 * n, *g0 = p.g0 + (size_t)b * c0;
    double *off = p.kkt0_ff + (size_t)b * s0;
    double *ofth = h > 0 ? p.kkt0_fth + (size_t)b * s0 * h : nullptr;
    double *x1th = X1 + n * (n + 1);      // theta columns next to x1v: one n x NA col-major right-hand-side block
    // G0 == I exactly (the initial-state constraint x0 = xhat of every MPC / DDP problem,
    // tests/gar/test_util.cpp:62-64)?  Then the second block row gives x0 = -g0 + mudyn lbd0 and
    // the first one collapses to (I + mudyn P0) lbd0 = P0 g0 - p0: an n x n SPD-type system of
    // the same form as Lambda above instead of the indefinite (2n)^2 Bunch-Kaufman problem.
    int isI = (c0 == n);
    if (isI)
      for (int e = tid; e < n * n; e += NT) {
        const int i = e % n, j = e / n;
        if (G0[e] != (i == j ? 1.0 : 0.0)) isI = 0;
      }
    isI = cta_all<NT, GRP>(isI);
    if (isI) {
      double fro = 0.0;
      for (int e = tid; e < n * n; e += NT) fro = fma(P[e], P[e], fro);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) fro += __shfl_xor_sync(0xffffffffu, fro, o);
      if (lane == 0) red[warp] = fro;
      gemv4<n, n, NT>(tid, [&](int i, int j) { return P[i + j * n]; }, [&](int j) { return g0[j]; },
                      [&](int i) { return -pv[i]; }, [&](int i, double v) { x1v[i] = v; });
      if constexpr (h > 0)
        for (int e = tid; e < n * h; e += NT) x1th[e] = -Pth[e];      // (I + mudyn P0) lbd0_th = -Vxt0
      cta_sync<NT, GRP>();
      fro = 0.0;
#pragma unroll
      for (int w = 0; w < NW; w++) fro += red[w];
      const double *lam = x1v;
      if (mudyn * sqrt(fro) <= 1e-9) {
        gemv4<n, n, NT>(tid, [&](int i, int j) { return -mudyn * P[i + j * n]; }, [&](int j) { return x1v[j]; },
                        [&](int i) { return x1v[i]; }, [&](int i, double v) { dinv[i] = v; });
        lam = dinv;
        if constexpr (h > 0) {
#pragma unroll 1
          for (int j = 0; j < h; j++)
            gemv4<n, n, NT>(tid, [&](int i, int k) { return -mudyn * P[i + k * n]; }, [&](int k) { return x1th[k + j * n]; },
                            [&](int i) { return x1th[i + j * n]; }, [&](int i, double v) { Pxt[i + j * n] = v; });
          cta_sync<NT, GRP>();
          for (int e = tid; e < n * h; e += NT) x1th[e] = Pxt[e];
        }
      } else {
        for (int e = tid; e < n * n; e += NT) {
          const int i = e % n, j = e / n;
          if (i >= j) X1[e] = (i == j ? 1.0 : 0.0) + mudyn * P[e];
        }
        cta_sync<NT, GRP>();
        if (!ldlt_solve_blocked<n, NA, NT, n, GRP>(tid, X1, RhsColMajor<n>{x1v}, Gs, dinv)) flags |= GARB200_FLAG_LLT_NOT_PD;
      }
      cta_sync<NT, GRP>();
      for (int i = tid; i < n; i += NT) { off[i] = fma(mudyn, lam[i], -g0[i]); off[n + i] = lam[i]; }
      if constexpr (h > 0) {
        // x0_th = mudyn lbd0_th ; thGrad = vt + Vxt^T x0 ; thHess = Vtt + Vxt^T x0_th         (:59-61)
        for (int e = tid; e < n * h; e += NT) {
          const int i = e / hh, j = e % hh;
          const double l = x1th[i + j * n];
          ofth[i * h + j] = mudyn * l;
          ofth[(n + i) * h + j] = l;
        }
        for (int e = tid; e < h * h + h; e += NT) {
          if (e < h * h) {
            const int a = e % hh, bb = e / hh;
            double acc = Vtt[e];
            for (int i = 0; i < n; i++) acc = fma(Pth[i + a * n], mudyn * x1th[i + bb * n], acc);
            p.thHess[(size_t)b * h * h + e] = acc;
          } else {
            const int a = e - h * h;
            double acc = vt[a];
            for (int i = 0; i < n; i++) acc = fma(Pth[i + a * n], fma(mudyn, lam[i], -g0[i]), acc);
            p.thGrad[(size_t)b * h + a] = acc;
          }
        }
      }
    } else {
      const int ld0 = s0 | 1;                        // odd leading dimension: conflict-free rows and columns
      double *K0 = sm;                               // overlays rec | X1 (size checked on the host)
      double *ff0 = Gs;
      double *wk0 = Gs + ((s0 + 1) & ~1);
      for (int e = tid; e < s0 * s0; e += NT) {
        const int i = e % s0, j = e / s0;
        double v;
        if (j < n) v = (i < n) ? P[i + j * n] : G0[(i - n) + j * c0];
        else v = (i == j) ? -mudyn : 0.0;
        K0[i + j * ld0] = v;
      }
      for (int i = tid; i < s0; i += NT) ff0[i] = (i < n) ? -pv[i] : -g0[i - n];
      cta_sync<NT, GRP>();
      if (!bk_factor_cta<NT, GRP>(tid, s0, K0, ld0, kpiv, ksub, wk0, kpiv + s0)) flags |= GARB200_FLAG_BK_SINGULAR;
      if (warp == 0) bk_solve_vec_warp<(2 * n + 31) / 32>(lane, s0, K0, ld0, kpiv, ksub, ff0, reinterpret_cast<int *>(wk0));
      cta_sync<NT, GRP>();
      for (int i = tid; i < s0; i += NT) off[i] = ff0[i];
      if constexpr (h > 0) {
        // theta columns one after the other through the same warp solve; X1 (free here? no: K0
        // overlays it) -> the columns live in Pxt|Vxn (2 n h >= s0 h doubles)
        double *fth0 = Pxt;
        static_assert(PL::Vxn == PL::Pxt + n * h, "Pxt and Vxn must be adjacent");
#pragma unroll 1
        for (int j = 0; j < h; j++) {
          cta_sync<NT, GRP>();
          for (int i = tid; i < s0; i += NT) ff0[i] = (i < n) ? -Pth[i + j * n] : 0.0;
          cta_sync<NT, GRP>();
          if (warp == 0) bk_solve_vec_warp<(2 * n + 31) / 32>(lane, s0, K0, ld0, kpiv, ksub, ff0, reinterpret_cast<int *>(wk0));
          cta_sync<NT, GRP>();
          for (int i = tid; i < s0; i += NT) { ofth[i * h + j] = ff0[i]; if (i < n) fth0[i + j * n] = ff0[i]; }
        }
        cta_sync<NT, GRP>();
        for (int e = tid; e < h * h + h; e += NT) {
          if (e < h * h) {
            const int a = e % hh, bb = e / hh;
            double acc = Vtt[e];
            for (int i = 0; i < n; i++) acc = fma(Pth[i + a * n], fth0[i + bb * n], acc);
            p.thHess[(size_t)b * h * h + e] = acc;
          } else {
            const int a = e - h * h;
            double acc = vt[a];
            for (int i = 0; i < n; i++) acc = fma(Pth[i + a * n], off[i], acc);
            p.thGrad[(size_t)b * h + a] = acc;
          }
        }
      }
    }
  }
  for (int e = tid; e < n * n; e += NT)
    if (!isfinite(P[e])) flags |= GARB200_FLAG_NONFINITE;
  if (flags) atomicOr(reinterpret_cast<int *>(p.flags) + b, flags);
  if (tid == 0) tma_store_wait_all();
  PH(15);
  if constexpr (PROF) {
    if (tid == 0 && p.phase_cycles)
#pragma unroll
      for (int i = 0; i < 16; i++) atomicAdd(reinterpret_cast<unsigned long long *>(p.phase_cycles) + i, (unsigned long long)ph_acc[i]);
  }
}

// ---- host side -----------------------------------------------------------------------------
template <int NX, int NU, int NC, int NTH, int NW, bool EG = false, int GRP = 1, bool EX = false>
static int launch_fast_t(const DevParams &p, cudaStream_t st) {
  using PL = FastPlan<NX, NU, NC, NTH, NTH, EG>;
  static_assert(!EG || PL::BtV + PL::KS + (PL::KS + 1) / 2 + 2 <= PL::kkt, "EG: BK scratch of the initial stage must fit in BtV");
  static_assert(PL::kstride >= 0, "");
  if (p.KL.stride != PL::kstride || p.GL.stride != PL::gstride || p.KL.E != PL::oE || p.KL.f != PL::of)
    return -3;   // layout mismatch: never expected, fall back to the generic kernel
  const int s0 = NX + p.nc0;
  if ((size_t)s0 * (s0 | 1) > (size_t)PL::P) return -3;      // K0 overlay must end before the P block
  // pre-pass
  constexpr int WPB = 4;
  constexpr int MPB = WPB * GjShape<NX>::MPW;
  const long long total = (long long)p.batch * p.T;
  int launched = 0;
  if (total > 0 && !EX) {
    const bool force_regs = p.tune.einv_mode == 1;
    if constexpr (NX % 4 == 0 && NX > 16) {
      if (!force_regs) {
        constexpr size_t esm = (size_t)WPB * (GjDmma<NX>::SMEM * sizeof(double) + GjDmma<NX>::SINTS * sizeof(int));
        if (cudaFuncSetAttribute(einv_dmma_kernel<NX, WPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)esm) != cudaSuccess) return -2;
        einv_dmma_kernel<NX, WPB><<<(unsigned)((total + WPB - 1) / WPB), 32 * WPB, esm, st>>>(p);
      }
      else einv_kernel<NX, WPB><<<(unsigned)((total + MPB - 1) / MPB), 32 * WPB, 0, st>>>(p);
    } else {
      // small matrices: the register cap trades a few spilled bytes for 24 / 32 resident warps per SM
      const int minb = p.tune.einv_minb;
      const unsigned grid = (unsigned)((total + MPB - 1) / MPB);
      if (minb >= 8) einv_kernel<NX, WPB, 8><<<grid, 32 * WPB, 0, st>>>(p);
      else if (minb >= 6) einv_kernel<NX, WPB, 6><<<grid, 32 * WPB, 0, st>>>(p);
      else einv_kernel<NX, WPB, 1><<<grid, 32 * WPB, 0, st>>>(p);
    }
    if (cudaGetLastError() != cudaSuccess) return -2;
    launched++;
  }
  const size_t smem = (size_t)GRP * PL::total * sizeof(double);
  static_assert((size_t)GRP * PL::total * sizeof(double) <= 227u * 1024u, "lockstep groups exceed the shared memory of one SM");
  auto ks = p.phase_cycles ? sweep_kernel<NX, NU, NC, NTH, NW, true, false, EG, GRP, EX> : sweep_kernel<NX, NU, NC, NTH, NW, false, false, EG, GRP, EX>;
  if (cudaFuncSetAttribute(ks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -2;
  ks<<<(p.batch + GRP - 1) / GRP, 32 * NW * GRP, smem, st>>>(p);
  if (cudaGetLastError() != cudaSuccess) return -2;
  return launched + 1;
}

// Leg mode (ParallelRiccatiSolver::backward, parallel-solver.hxx:111-143): the E^{-1} pre-pass, then the
// coupled legs (theta = co-state of the next leg, NTH = NX) and the last leg (NTH = 0).
template <int NX, int NU, int NW>
static int launch_fast_legs(const DevParams &p, cudaStream_t st) {
  using PLc = FastPlan<NX, NU, 0, NX, 0>;
  using PLl = FastPlan<NX, NU, 0, 0, 0>;
  if (p.KL.stride != PLc::kstride || p.GL.stride != PLc::gstride || p.KL.E != PLc::oE || p.KL.f != PLc::of) return -3;
  if ((size_t)PLc::total * sizeof(double) > 227u * 1024u) return -3;
  constexpr int WPB = 4;
  const long long total = (long long)p.batch * p.T;
  constexpr size_t esm = (size_t)WPB * (GjDmma<NX>::SMEM * sizeof(double) + GjDmma<NX>::SINTS * sizeof(int));
  if (cudaFuncSetAttribute(einv_dmma_kernel<NX, WPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)esm) != cudaSuccess) return -2;
  einv_dmma_kernel<NX, WPB><<<(unsigned)((total + WPB - 1) / WPB), 32 * WPB, esm, st>>>(p);
  if (cudaGetLastError() != cudaSuccess) return -2;
  auto kc = sweep_kernel<NX, NU, 0, NX, NW, false, true>;
  auto kl = sweep_kernel<NX, NU, 0, 0, NW, false, true>;
  const size_t smc = (size_t)PLc::total * sizeof(double), sml = (size_t)PLl::total * sizeof(double);
  if (cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc) != cudaSuccess) return -2;
  if (cudaFuncSetAttribute(kl, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sml) != cudaSuccess) return -2;
  // The last leg is independent of the coupled legs (disjoint knots, gains and value slots): it runs on
  // the solver's side stream (created in garb200_create) between two events, beside the coupled legs.
  struct Side { cudaStream_t s; cudaEvent_t e0, e1; bool ok; };
  Side side_{static_cast<cudaStream_t>(p.side_stream), static_cast<cudaEvent_t>(p.side_e0), static_cast<cudaEvent_t>(p.side_e1), p.side_stream != nullptr};
  Side *sd = &side_;
  const bool fork = sd && sd->ok && p.batch * (p.num_legs - 1) < 4 * 148;   // only worth it while the GPU is not full
  if (fork) {
    if (cudaEventRecord(sd->e0, st) != cudaSuccess || cudaStreamWaitEvent(sd->s, sd->e0, 0) != cudaSuccess) return -2;
    kl<<<p.batch, 32 * NW, sml, sd->s>>>(p);
    if (cudaGetLastError() != cudaSuccess || cudaEventRecord(sd->e1, sd->s) != cudaSuccess) return -2;
  }
  kc<<<p.batch * (p.num_legs - 1), 32 * NW, smc, st>>>(p);
  if (cudaGetLastError() != cudaSuccess) return -2;
  if (fork) {
    if (cudaStreamWaitEvent(st, sd->e1, 0) != cudaSuccess) return -2;
  } else {
    kl<<<p.batch, 32 * NW, sml, st>>>(p);
    if (cudaGetLastError() != cudaSuccess) return -2;
  }
  return 3;
}

/// Returns the number of kernels launched, -3 if no specialisation matches (caller falls back
/// to the generic kernel), < 0 otherwise on error.
int launch_backward_fast(const DevParams &p, const int *h_nu_t, const int *h_nc_t, cudaStream_t st) {
  if (p.einv == nullptr || p.T < 1) return -3;
  if (h_nu_t[p.T] != 0 || h_nc_t[p.T] != 0) return -3;
  for (int t = 0; t < p.T; t++)
    if (h_nu_t[t] != p.nu || h_nc_t[t] != p.nc) return -3;
  if (p.num_legs >= 2) {
    if (p.tune.legs_generic || p.store_all_values) return -3;
    if (p.nx == 36 && p.nu == 12 && p.nc == 0 && p.nth == 0) return launch_fast_legs<36, 12, 5>(p, st);
    return -3;
  }
  if (p.explicit_dyn) {       // every E_t == -I (verified on the device by detect_explicit_kernel, or promised by the caller)
    if (p.nx == 12 && p.nu == 4 && p.nc == 0 && p.nth == 0) return launch_fast_t<12, 4, 0, 0, 1, false, 1, true>(p, st);
    if (p.nx == 36 && p.nu == 12 && p.nc == 0 && p.nth == 0) return launch_fast_t<36, 12, 0, 0, 5, false, 1, true>(p, st);
    if (p.nx == 36 && p.nu == 12 && p.nc == 6 && p.nth == 4) return launch_fast_t<36, 12, 6, 4, 5, true, 1, true>(p, st);
    if (p.nx == 36 && p.nu == 12 && p.nc == 12 && p.nth == 8) return launch_fast_t<36, 12, 12, 8, 5, true, 1, true>(p, st);
  }
  if (p.nx == 12 && p.nu == 4 && p.nc == 0 && p.nth == 0) return launch_fast_t<12, 4, 0, 0, 1>(p, st);
  if (p.nx == 36 && p.nu == 12 && p.nc == 0 && p.nth == 0) {
    // lockstep groups of 3 problems per CTA are opt-in: measured SLOWER than 3 independent CTAs per SM
    // (c1b 28.4 vs 22.5 ms backward: the alignment barrier costs more than the chains gain)
    return p.tune.group >= 2 ? launch_fast_t<36, 12, 0, 0, 5, false, 3>(p, st) : launch_fast_t<36, 12, 0, 0, 5>(p, st);
  }
  if (p.nx == 36 && p.nu == 12 && p.nc == 6 && p.nth == 4) {   // C3: E^{-1} from L2 -> 75 KB of shared memory, 3 CTAs per SM
    return p.tune.c3_eg ? launch_fast_t<36, 12, 6, 4, 5, true>(p, st) : launch_fast_t<36, 12, 6, 4, 5, false>(p, st);
  }
  // the widest constrained / parametric instantiation: the covering shape of everything with nc <= 12, nth <= 8 (pad.cu)
  if (p.nx == 36 && p.nu == 12 && p.nc == 12 && p.nth == 8) return launch_fast_t<36, 12, 12, 8, 5, true>(p, st);
  if (big_path_available(p.nx, p.nu, p.nc, p.nth)) return launch_backward_big(p, st);                     // C5
  return -3;
}

bool fast_path_available(int nx, int nu, int nc, int nth, int num_legs) {
  if (num_legs >= 2) return nx == 36 && nu == 12 && nc == 0 && nth == 0;
  return (nx == 12 && nu == 4 && nc == 0 && nth == 0) || (nx == 36 && nu == 12 && nc == 0 && nth == 0) ||
         (nx == 36 && nu == 12 && nc == 6 && nth == 4) || (nx == 36 && nu == 12 && nc == 12 && nth == 8) ||
         big_path_available(nx, nu, nc, nth);
}

} // namespace garb200
