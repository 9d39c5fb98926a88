// backward.cu — batched proximal Riccati backward sweep (sm_100a).
//
// One thread group (a CTA of NT threads) owns one LQ problem — or, in parallel-in-time
// mode, one LEG of one problem — and walks its knots from the last to the first:
//   terminal_solve   ProximalRiccatiKernel::terminalSolve     riccati-kernel.hxx:90-150
//   stage_solve      ProximalRiccatiKernel::stageKernelSolve  riccati-kernel.hxx:167-300
//   (loop)           ProximalRiccatiKernel::backwardImpl      riccati-kernel.hxx:64-88
//   initial_solve    ProximalRiccatiSolver::backward          proximal-riccati.hxx:44-62
// The knot record is staged into shared memory with ONE 1-D TMA bulk copy
// (cp.async.bulk + mbarrier), every intermediate lives in shared memory, and the gains
// record of the knot is assembled in shared memory in its final global layout and leaves
// with ONE bulk store.  HBM traffic per knot = knot record in + gains record out.
#include "device_linalg.cuh"
#include "kernels.cuh"

namespace garb200 {

// ---- shared-memory plan (offsets in doubles) ------------------------------------------
struct SmemPlan {
  int rec, X1, X3, P, pv, vx, tv, BtV, kkt, ksub, G, Vxt, Vtt, vt, ints, bar, total;
};

__host__ __device__ inline SmemPlan make_plan(const DevParams &p) {
  const int n = p.nx, m = p.nu, c = p.nc;
  const int h = p.num_legs >= 2 ? p.hleg : p.nth;
  const int s = m + c;
  SmemPlan L;
  int o = 0;
  auto take = [&](int cnt) { int at = o; o = (o + cnt + 1) & ~1; return at; };
  // rec | X1 | X3 are contiguous: the initial-stage KKT matrix (nx+nc0)^2 overlays them
  L.rec = take(p.KL.stride);
  L.X1 = take(n * (n > h ? n : h));
  L.X3 = take(n * n);
  L.P = take(n * n);
  L.pv = take(n);
  L.vx = take(n);
  L.tv = take(n > h ? n : h);
  L.BtV = take(m * n);
  L.kkt = take(s * s);
  L.ksub = take((s > n + p.nc0 ? s : n + p.nc0) + 2);
  L.G = take(p.GL.stride);
  L.Vxt = take(n * h);
  L.Vtt = take(h * h);
  L.vt = take(h);
  // ints: BK pivots (max(s, nx+nc0)) + LU scratch 3n   (2 ints per double slot)
  const int nints = (s > n + p.nc0 ? s : n + p.nc0) + 3 * n + 4;
  L.ints = take((nints + 1) / 2);
  L.bar = take(2);
  L.total = o;
  return L;
}

/// Strided read-only view of the parameter block of a knot (or of the implicit one a
/// leg-end knot carries: Gx = A^T, Gu = B^T, Gth = -mu I, gamma = f; parallel-solver.hpp:43-51).
struct ParamBlock {
  const double *Gx; int gx_rs, gx_cs;   // nx x h
  const double *Gu; int gu_rs, gu_cs;   // nu x h
  const double *Gth;                    // h x h col-major, or nullptr
  double gth_diag;                      // used when Gth == nullptr: Gth = gth_diag * I
  const double *gamma;                  // h
};

template <class G> struct Ctx {
  G g;
  int n, h;            // state dim, effective parameter dim of this sweep
  double mudyn, mueq;
  double *rec, *X1, *X3, *P, *pv, *vx, *tv, *BtV, *kkt, *ksub, *Gs, *Vxt, *Vtt, *vt;
  int *kpiv, *iw;
  garb200_knot_layout KL;
  garb200_gain_layout GL;
  int flags;
};

// BK factor (warp 0) + column-parallel solves of the (1 + n + h) right-hand sides that sit
// in the gains staging record: ff[0:s], fb rows 0..s-1, fth rows 0..s-1.
template <class G>
__device__ __forceinline__ void reduced_kkt_solve(Ctx<G> &cx, int s, bool with_fb) {
  const G &g = cx.g;
  const int n = cx.n, h = cx.h;
  g.sync();
  if (g.warp() == 0) {
    const bool ok = bk_factor_warp(g.lane(), s, cx.kkt, s, cx.kpiv, cx.ksub);
    if (!ok) cx.flags |= GARB200_FLAG_BK_SINGULAR;
  }
  g.sync();
  double *ff = cx.Gs + cx.GL.ff, *fb = cx.Gs + cx.GL.fb, *fth = cx.Gs + cx.GL.fth;
  const int ncols = 1 + (with_fb ? n : 0) + h;
  for (int col = g.tid; col < ncols; col += G::NT) {
    double *xp; int xs;
    if (col == 0) { xp = ff; xs = 1; }
    else if (with_fb && col <= n) { xp = fb + (col - 1); xs = n; }
    else { xp = fth + (col - 1 - (with_fb ? n : 0)); xs = h; }
    bk_solve_column(s, cx.kkt, s, cx.kpiv, cx.ksub, xp, xs);
  }
  g.sync();
}

// assemble the lower triangle of [Rm D^T; D -mueq I] into cx.kkt (Rm already in place)
template <class G>
__device__ __forceinline__ void fill_kkt_constraint_rows(Ctx<G> &cx, int m, int c,
                                                         const double *D) {
  const int s = m + c;
  for (int e = cx.g.tid; e < c * s; e += G::NT) {
    const int i = e % c, j = e / c;  // row m+i, col j (j < s)
    double v;
    if (j < m) v = D[i + j * c];
    else v = (j - m == i) ? -cx.mueq : 0.0;
    cx.kkt[(m + i) + j * s] = v;
  }
}

// ---- terminal knot --------------------------------------------------------------------
template <int TM, int TN, class G>
__device__ __forceinline__ void terminal_solve(Ctx<G> &cx, int m, int c, const ParamBlock &pb) {
  const G &g = cx.g;
  const int n = cx.n, h = cx.h, s = m + c;
  const garb200_knot_layout &KL = cx.KL;
  double *rec = cx.rec;
  double *Q = rec + KL.Q, *S = rec + KL.S, *R = rec + KL.R, *q = rec + KL.q, *r = rec + KL.r;
  double *C = rec + KL.C, *D = rec + KL.D, *d = rec + KL.d;
  double *ff = cx.Gs + cx.GL.ff, *fb = cx.Gs + cx.GL.fb, *fth = cx.Gs + cx.GL.fth;
  for (int e = g.tid; e < cx.GL.stride; e += G::NT) cx.Gs[e] = 0.0;
  g.sync();
  if (m == 0) {                                                    // :105-108
    for (int e = g.tid; e < c * n; e += G::NT) { const int i = e % c, j = e / c; fb[i * n + j] = C[i + j * c] / cx.mueq; }
    for (int i = g.tid; i < c; i += G::NT) ff[i] = d[i] / cx.mueq;
    g.sync();
  } else {                                                         // :110-131
    for (int e = g.tid; e < m * m; e += G::NT) { const int i = e % m, j = e / m; cx.kkt[i + j * s] = R[i + j * m]; }
    fill_kkt_constraint_rows(cx, m, c, D);
    for (int i = g.tid; i < m; i += G::NT) ff[i] = -r[i];
    for (int i = g.tid; i < c; i += G::NT) ff[m + i] = -d[i];
    copy_mat(g, m, n, S, n, 1, -1.0, fb, n, 1);                    // K rows = -S^T
    copy_mat(g, c, n, C, 1, c, -1.0, fb + m * n, n, 1);            // Z rows = -C
    if (h > 0 && pb.Gu) copy_mat(g, m, h, pb.Gu, pb.gu_rs, pb.gu_cs, -1.0, fth, h, 1);
    reduced_kkt_solve(cx, s, true);
  }
  // Pmat = Q + C^T Z (+ S K) ; pvec = q + C^T zff (+ S kff)           (:134-140)
  gemm<TM, TN>(g, n, n, c, 1.0, C, c, 1, fb + m * n, n, 1, 1.0, Q, 1, n, cx.P, 1, n);
  if (m > 0) gemm<TM, TN>(g, n, n, m, 1.0, S, 1, n, fb, n, 1, 1.0, cx.P, 1, n, cx.P, 1, n);
  gemv(g, n, c, 1.0, C, c, 1, ff + m, 1, 1.0, q, 1, cx.pv, 1);
  if (m > 0) gemv(g, n, m, 1.0, S, 1, n, ff, 1, 1.0, cx.pv, 1, cx.pv, 1);
  if (h > 0) {                                                     // :142-149
    // Vxt = Gx + K^T Gu ; Vtt = Gth + Gu^T Kth ; vt = gamma + Gu^T kff
    for (int e = g.tid; e < n * h; e += G::NT) {
      const int i = e % n, j = e / n;
      cx.Vxt[e] = pb.Gx ? pb.Gx[i * pb.gx_rs + j * pb.gx_cs] : 0.0;
    }
    for (int e = g.tid; e < h * h; e += G::NT) {
      const int i = e % h, j = e / h;
      cx.Vtt[e] = pb.Gth ? pb.Gth[e] : (i == j ? pb.gth_diag : 0.0);
    }
    for (int i = g.tid; i < h; i += G::NT) cx.vt[i] = pb.gamma ? pb.gamma[i] : 0.0;
    g.sync();
    if (m > 0 && pb.Gu) {
      gemm<TM, TN>(g, n, h, m, 1.0, fb, 1, n, pb.Gu, pb.gu_rs, pb.gu_cs, 1.0, cx.Vxt, 1, n, cx.Vxt, 1, n);
      gemm<TM, TN>(g, h, h, m, 1.0, pb.Gu, pb.gu_cs, pb.gu_rs, fth, h, 1, 1.0, cx.Vtt, 1, h, cx.Vtt, 1, h);
      gemv(g, h, m, 1.0, pb.Gu, pb.gu_cs, pb.gu_rs, ff, 1, 1.0, cx.vt, 1, cx.vt, 1);
    }
  }
  g.sync();
}

// ---- one non-terminal knot ------------------------------------------------------------
// On entry cx.P/pv (+Vxt/Vtt/vt) hold the value function of knot t+1 (`vn`); on exit they
// hold knot t's (`vc`) and cx.Gs holds knot t's gains record.  If vnext != nullptr the
// intermediate Vxx / vx (what the reference files under datas[t+1].vm) are written there.
template <int TM, int TN, class G>
__device__ __forceinline__ void stage_solve(Ctx<G> &cx, int m, int c, const ParamBlock &pb,
                                            double *vnext_Vxx, double *vnext_vx) {
  const G &g = cx.g;
  const int n = cx.n, h = cx.h, s = m + c;
  const double mudyn = cx.mudyn;
  const garb200_knot_layout &KL = cx.KL;
  double *rec = cx.rec;
  double *Q = rec + KL.Q, *S = rec + KL.S, *R = rec + KL.R, *q = rec + KL.q, *r = rec + KL.r;
  double *A = rec + KL.A, *B = rec + KL.B, *E = rec + KL.E, *f = rec + KL.f;
  double *C = rec + KL.C, *D = rec + KL.D, *d = rec + KL.d;
  double *X1 = cx.X1, *Einv = cx.X3, *P = cx.P, *pv = cx.pv, *vx = cx.vx, *tv = cx.tv;
  double *BtV = cx.BtV;
  double *ff = cx.Gs + cx.GL.ff, *fb = cx.Gs + cx.GL.fb, *fth = cx.Gs + cx.GL.fth;
  const int rK = 0, rZ = m, rL = m + c, rA = m + c + n;
  double *Lth = fth + rL * h, *Kth = fth + rK * h, *Yth = fth + rA * h;

  // (1) Einv = PartialPivLU(E).solve(I)                                  :172-174
  if (!lu_inverse(g, n, E, X1, Einv, cx.iw)) cx.flags |= GARB200_FLAG_LU_SINGULAR;
  // (2) ptilde = -Einv^T p ; EinvP = Einv^T P ; Pxttilde = -Einv^T Vxt   :175-178, :261-262
  gemv(g, n, n, -1.0, Einv, n, 1, pv, 1, 0.0, nullptr, 0, vx, 1);
  gemm<TM, TN>(g, n, n, n, 1.0, Einv, n, 1, P, 1, n, 0.0, nullptr, 0, 0, X1, 1, n);
  if (h > 0) {
    for (int e = g.tid; e < cx.GL.nrows * h; e += G::NT) fth[e] = 0.0;
    g.sync();
    gemm<TM, TN>(g, n, h, n, -1.0, Einv, n, 1, cx.Vxt, 1, n, 0.0, nullptr, 0, 0, Lth, h, 1);
  }
  g.sync();
  // (3) Ptilde = EinvP * Einv, lower-symmetrised                         :179-180
  gemm<TM, TN>(g, n, n, n, 1.0, X1, 1, n, Einv, 1, n, 0.0, nullptr, 0, 0, P, 1, n);
  g.sync();
  mirror_lower(g, n, P, n);
  g.sync();
  // (4) Lambda = I + mudyn Ptilde ; Vxx = Lambda^{-1} Ptilde ; vx = Lambda^{-1}(ptilde + Ptilde f)
  gemv(g, n, n, 1.0, P, 1, n, f, 1, 1.0, vx, 1, vx, 1);              // :187
  for (int e = g.tid; e < n * n; e += G::NT) X1[e] = ((e % n) == (e / n) ? 1.0 : 0.0) + mudyn * P[e];
  g.sync();
  {
    RhsView r0{P, 1, n, n}, r1{vx, 1, 0, 1}, r2{Lth, h, 1, h};        // :185-190, :265
    if (!spd_solve(g, n, X1, n, r0, r1, r2)) cx.flags |= GARB200_FLAG_LLT_NOT_PD;
  }
  mirror_lower(g, n, P, n);                                            // :190
  g.sync();
  if (vnext_Vxx) {
    for (int e = g.tid; e < n * n; e += G::NT) vnext_Vxx[e] = P[e];
    for (int i = g.tid; i < n; i += G::NT) vnext_vx[i] = vx[i];
  }
  // (5) AtV = A^T Vxx -> X1 ; BtV = B^T Vxx                              :192-193
  gemm<TM, TN>(g, n, n, n, 1.0, A, n, 1, P, 1, n, 0.0, nullptr, 0, 0, X1, 1, n);
  gemm<TM, TN>(g, m, n, n, 1.0, B, n, 1, P, 1, n, 0.0, nullptr, 0, 0, BtV, 1, m);
  g.sync();
  // (6) Qhat, Shat, Rhat, qhat, rhat (in place in the staged record)     :195-199
  gemm<TM, TN>(g, n, n, n, 1.0, X1, 1, n, A, 1, n, 1.0, Q, 1, n, Q, 1, n);
  gemm<TM, TN>(g, n, m, n, 1.0, X1, 1, n, B, 1, n, 1.0, S, 1, n, S, 1, n);
  gemm<TM, TN>(g, m, m, n, 1.0, BtV, 1, m, B, 1, n, 1.0, R, 1, m, cx.kkt, 1, s);
  gemv(g, n, n, 1.0, A, n, 1, vx, 1, 1.0, q, 1, q, 1);
  gemv(g, m, n, -1.0, B, n, 1, vx, 1, -1.0, r, 1, ff + rK, 1);      // kff rhs = -rhat   :216
  fill_kkt_constraint_rows(cx, m, c, D);                               // :202-207
  for (int i = g.tid; i < c; i += G::NT) ff[rZ + i] = -d[i];        // :217
  copy_mat(g, c, n, C, 1, c, -1.0, fb + rZ * n, n, 1);                 // Z rhs = -C        :224
  if (h > 0) {                                                         // Guhat, :268-272
    gemm<TM, TN>(g, m, h, n, -1.0, B, n, 1, Lth, h, 1, -1.0, pb.Gu, pb.gu_rs, pb.gu_cs, Kth, h, 1);
  }
  g.sync();
  copy_mat(g, m, n, S, n, 1, -1.0, fb + rK * n, n, 1);                 // K rhs = -Shat^T   :223
  reduced_kkt_solve(cx, s, true);                                      // :208, :227-228, :274
  // (8) closed-loop feed-forward                                         :231-236
  gemv(g, n, m, 1.0, BtV, m, 1, ff + rK, 1, 1.0, vx, 1, ff + rL, 1); // lff
  gemv(g, n, m, 1.0, B, 1, n, ff + rK, 1, 1.0, f, 1, tv, 1);         // f + B kff
  // (9) L = Vxx A + BtV^T K                                              :238-239
  gemm<TM, TN>(g, n, n, n, 1.0, P, 1, n, A, 1, n, 0.0, nullptr, 0, 0, fb + rL * n, n, 1);
  gemm<TM, TN>(g, n, n, m, 1.0, BtV, m, 1, fb + rK * n, n, 1, 1.0, fb + rL * n, n, 1, fb + rL * n, n, 1);
  if (h > 0)                                                           // Lth += BtV^T Kth  :277
    gemm<TM, TN>(g, n, h, m, 1.0, BtV, m, 1, Kth, h, 1, 1.0, Lth, h, 1, Lth, h, 1);
  g.sync();
  for (int i = g.tid; i < n; i += G::NT) tv[i] -= mudyn * ff[rL + i];
  // (10) A_pre = A + B K - mudyn L (in place) ; Afb = -Einv A_pre         :241-245
  gemm<TM, TN>(g, n, n, m, 1.0, B, 1, n, fb + rK * n, n, 1, 1.0, A, 1, n, A, 1, n);
  g.sync();
  for (int e = g.tid; e < n * n; e += G::NT) { const int i = e % n, j = e / n; A[e] -= mudyn * fb[(rL + i) * n + j]; }
  gemv(g, n, n, -1.0, Einv, 1, n, tv, 1, 0.0, nullptr, 0, ff + rA, 1); // yff
  g.sync();
  gemm<TM, TN>(g, n, n, n, -1.0, Einv, 1, n, A, 1, n, 0.0, nullptr, 0, 0, fb + rA * n, n, 1);
  // (11) Pmat = Qhat + Shat K + C^T Z ; pvec = qhat + Shat kff + C^T zff   :249-250
  gemm<TM, TN>(g, n, n, m, 1.0, S, 1, n, fb + rK * n, n, 1, 1.0, Q, 1, n, P, 1, n);
  if (c > 0) gemm<TM, TN>(g, n, n, c, 1.0, C, c, 1, fb + rZ * n, n, 1, 1.0, P, 1, n, P, 1, n);
  gemv(g, n, m, 1.0, S, 1, n, ff + rK, 1, 1.0, q, 1, pv, 1);
  if (c > 0) gemv(g, n, c, 1.0, C, c, 1, ff + rZ, 1, 1.0, pv, 1, pv, 1);
  if (h > 0) {                                                         // :279-298
    // Yth_pre = B Kth - mudyn Lth -> X1 ; Yth = -Einv Yth_pre
    gemm<TM, TN>(g, n, h, m, 1.0, B, 1, n, Kth, h, 1, -mudyn, Lth, h, 1, X1, 1, n);
    g.sync();
    gemm<TM, TN>(g, n, h, n, -1.0, Einv, 1, n, X1, 1, n, 0.0, nullptr, 0, 0, Yth, h, 1);
    // vt = vn.vt + gamma + Gu^T kff + Vxt^T yff
    for (int i = g.tid; i < h; i += G::NT) {
      double acc = cx.vt[i] + (pb.gamma ? pb.gamma[i] : 0.0);
      if (pb.Gu) for (int k = 0; k < m; k++) acc = fma(pb.Gu[k * pb.gu_rs + i * pb.gu_cs], ff[rK + k], acc);
      for (int k = 0; k < n; k++) acc = fma(cx.Vxt[k + i * n], ff[rA + k], acc);
      cx.vt[i] = acc;
    }
    g.sync();
    // Vtt += Gth + Gu^T Kth + Vxt^T Yth   (uses the OLD Vxt)
    if (pb.Gu) gemm<TM, TN>(g, h, h, m, 1.0, pb.Gu, pb.gu_cs, pb.gu_rs, Kth, h, 1, 1.0, cx.Vtt, 1, h, cx.Vtt, 1, h);
    gemm<TM, TN>(g, h, h, n, 1.0, cx.Vxt, n, 1, Yth, h, 1, 1.0, cx.Vtt, 1, h, cx.Vtt, 1, h);
    // Vxt_new = Gx + K^T Gu + Afb^T Vxt -> X1 (Yth_pre is consumed)
    gemm<TM, TN>(g, n, h, n, 1.0, fb + rA * n, 1, n, cx.Vxt, 1, n, pb.Gx ? 1.0 : 0.0, pb.Gx, pb.gx_rs, pb.gx_cs, X1, 1, n);
    if (pb.Gu) gemm<TM, TN>(g, n, h, m, 1.0, fb + rK * n, 1, n, pb.Gu, pb.gu_rs, pb.gu_cs, 1.0, X1, 1, n, X1, 1, n);
    g.sync();
    for (int e = g.tid; e < n * h; e += G::NT) cx.Vxt[e] = X1[e];
    for (int e = g.tid; e < h * h; e += G::NT) {
      const int i = e % h, j = e / h;
      cx.Vtt[e] += pb.Gth ? pb.Gth[e] : (i == j ? pb.gth_diag : 0.0);
    }
  }
  g.sync();
}

// ---- initial stage: BK of [Pmat0 G0^T; G0 -mudyn I]  (proximal-riccati.hxx:44-62) -------
template <class G>
__device__ __forceinline__ void initial_solve(Ctx<G> &cx, const DevParams &p, int b) {
  const G &g = cx.g;
  const int n = cx.n, h = cx.h, c0 = p.nc0, s0 = n + c0;
  double *K0 = cx.rec;                       // overlays rec | X1 | X3 (>= 4 n^2 doubles)
  double *ff0 = cx.Gs, *fth0 = cx.Gs + ((s0 + 1) & ~1);
  const double *G0 = p.G0 + (size_t)b * c0 * n, *g0 = p.g0 + (size_t)b * c0;
  for (int e = g.tid; e < s0 * s0; e += G::NT) {
    const int i = e % s0, j = e / s0;
    double v;
    if (j < n) v = (i < n) ? cx.P[i + j * n] : G0[(i - n) + j * c0];
    else v = (i == j) ? -cx.mudyn : 0.0;
    K0[e] = v;
  }
  for (int i = g.tid; i < s0; i += G::NT) ff0[i] = (i < n) ? -cx.pv[i] : -g0[i - n];
  for (int e = g.tid; e < s0 * h; e += G::NT) {
    const int i = e / h, j = e % h;
    fth0[e] = (i < n) ? -cx.Vxt[i + j * n] : 0.0;
  }
  g.sync();
  if (g.warp() == 0) {
    if (!bk_factor_warp(g.lane(), s0, K0, s0, cx.kpiv, cx.ksub)) cx.flags |= GARB200_FLAG_BK_SINGULAR;
  }
  g.sync();
  for (int col = g.tid; col < 1 + h; col += G::NT) {
    if (col == 0) bk_solve_column(s0, K0, s0, cx.kpiv, cx.ksub, ff0, 1);
    else bk_solve_column(s0, K0, s0, cx.kpiv, cx.ksub, fth0 + (col - 1), h);
  }
  g.sync();
  double *off = p.kkt0_ff + (size_t)b * s0, *ofth = p.kkt0_fth + (size_t)b * s0 * h;
  for (int i = g.tid; i < s0; i += G::NT) off[i] = ff0[i];
  for (int e = g.tid; e < s0 * h; e += G::NT) ofth[e] = fth0[e];
  // thGrad = vt + Vxt^T x0 ; thHess = Vtt + Vxt^T fth_x                  (:59-61)
  for (int a = g.tid; a < h; a += G::NT) {
    double acc = cx.vt[a];
    for (int i = 0; i < n; i++) acc = fma(cx.Vxt[i + a * n], ff0[i], acc);
    p.thGrad[(size_t)b * h + a] = acc;
  }
  for (int e = g.tid; e < h * h; e += G::NT) {
    const int a = e % h, bb = e / h;
    double acc = cx.Vtt[e];
    for (int i = 0; i < n; i++) acc = fma(cx.Vxt[i + a * n], fth0[i * h + bb], acc);
    p.thHess[(size_t)b * h * h + e] = acc;
  }
  g.sync();
}

// ---- the sweep kernel ------------------------------------------------------------------
template <int NT, int TM, int TN>
__global__ void __launch_bounds__(NT) backward_kernel(const DevParams p) {
  extern __shared__ __align__(16) double sm[];
  using G = Group<NT>;
  const SmemPlan pl = make_plan(p);
  const int nlegs = p.num_legs >= 2 ? p.num_legs : 1;
  const int b = blockIdx.x / nlegs, leg = blockIdx.x % nlegs;
  int beg = 0, end = p.T + 1;
  if (nlegs > 1) get_work(p.T, leg, nlegs, beg, end);
  const bool legmode = nlegs > 1;
  const bool coupled_leg = legmode && leg < nlegs - 1;  // leg ends on a co-state cut

  Ctx<G> cx;
  cx.g.tid = threadIdx.x; cx.g.bar_id = 0;
  cx.n = p.nx;
  cx.h = legmode ? (coupled_leg ? p.hleg : 0) : p.nth;
  cx.mudyn = p.mudyn; cx.mueq = p.mueq;
  cx.rec = sm + pl.rec; cx.X1 = sm + pl.X1; cx.X3 = sm + pl.X3; cx.P = sm + pl.P;
  cx.pv = sm + pl.pv; cx.vx = sm + pl.vx; cx.tv = sm + pl.tv; cx.BtV = sm + pl.BtV;
  cx.kkt = sm + pl.kkt; cx.ksub = sm + pl.ksub; cx.Gs = sm + pl.G;
  cx.Vxt = sm + pl.Vxt; cx.Vtt = sm + pl.Vtt; cx.vt = sm + pl.vt;
  cx.kpiv = reinterpret_cast<int *>(sm + pl.ints);
  cx.iw = cx.kpiv + ((p.nu + p.nc) > (p.nx + p.nc0) ? (p.nu + p.nc) : (p.nx + p.nc0)) + 1;
  cx.KL = p.KL; cx.GL = p.GL; cx.flags = 0;
  uint64_t *bar = reinterpret_cast<uint64_t *>(sm + pl.bar);
  const G &g = cx.g;
  const int n = p.nx, h = cx.h;

  if (g.tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  g.sync();
  uint32_t phase = 0;
  const uint32_t rec_bytes = (uint32_t)p.KL.stride * 8u;
  const size_t pbase = (size_t)b * (p.T + 1);

  for (int t = end - 1; t >= beg; t--) {
    const int pt = phys_t(p, t);
    if (g.tid == 0) {
      fence_proxy_async();
      mbar_expect_tx(bar, rec_bytes);
      tma_load_1d(cx.rec, p.knots + (pbase + pt) * p.KL.stride, rec_bytes, bar);
      tma_store_wait_read();  // previous gains record has left shared memory
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    g.sync();
    const int m = p.nu_t[t], c = p.nc_t[t];
    ParamBlock pb;
    const garb200_knot_layout &KL = p.KL;
    if (!legmode) {
      pb.Gx = h ? cx.rec + KL.Gx : nullptr; pb.gx_rs = 1; pb.gx_cs = n;
      pb.Gu = h ? cx.rec + KL.Gu : nullptr; pb.gu_rs = 1; pb.gu_cs = m;
      pb.Gth = h ? cx.rec + KL.Gth : nullptr; pb.gth_diag = 0.0;
      pb.gamma = h ? cx.rec + KL.gamma : nullptr;
    } else if (coupled_leg && t == end - 1) {   // setupKnot, parallel-solver.hpp:43-51
      pb.Gx = cx.rec + KL.A; pb.gx_rs = n; pb.gx_cs = 1;   // Gx = A^T
      pb.Gu = cx.rec + KL.B; pb.gu_rs = n; pb.gu_cs = 1;   // Gu = B^T
      pb.Gth = nullptr; pb.gth_diag = -p.mudyn;
      pb.gamma = cx.rec + KL.f;
    } else {                                    // zero parameter block (allocateLeg :51-61)
      pb.Gx = nullptr; pb.Gu = nullptr; pb.Gth = nullptr; pb.gth_diag = 0.0; pb.gamma = nullptr;
      pb.gx_rs = pb.gx_cs = pb.gu_rs = pb.gu_cs = 0;
    }
    double *vrec_next = nullptr;
    if (p.store_all_values && t + 1 <= p.T)
      vrec_next = p.values + ((size_t)b * p.nslots + (t + 1)) * p.VL.stride;
    if (t == end - 1) terminal_solve<TM, TN>(cx, m, c, pb);
    else stage_solve<TM, TN>(cx, m, c, pb, vrec_next ? vrec_next + p.VL.Vxx : nullptr,
                             vrec_next ? vrec_next + p.VL.vx : nullptr);
    // gains record -> global (one bulk store)
    fence_proxy_async();
    g.sync();
    if (g.tid == 0) {
      tma_store_1d(p.gains + (pbase + pt) * p.GL.stride, cx.Gs, (uint32_t)p.GL.stride * 8u);
      tma_store_commit();
    }
    // value function of this knot (what the algorithm / callers consume)
    const int slot = p.store_all_values ? t : (t == beg ? leg : -1);
    if (slot >= 0) {
      double *v = p.values + ((size_t)b * p.nslots + slot) * p.VL.stride;
      for (int e = g.tid; e < n * n; e += NT) v[p.VL.Pmat + e] = cx.P[e];
      for (int i = g.tid; i < n; i += NT) v[p.VL.pvec + i] = cx.pv[i];
      for (int e = g.tid; e < n * h; e += NT) v[p.VL.Vxt + e] = cx.Vxt[e];
      for (int e = g.tid; e < h * h; e += NT) v[p.VL.Vtt + e] = cx.Vtt[e];
      for (int i = g.tid; i < h; i += NT) v[p.VL.vt + i] = cx.vt[i];
      if (t == 0 && !legmode && p.store_all_values) {   // proximal-riccati.hxx:39-42
        for (int e = g.tid; e < n * n; e += NT) v[p.VL.Vxx + e] = cx.P[e];
        for (int i = g.tid; i < n; i += NT) v[p.VL.vx + i] = cx.pv[i];
      }
    }
    fence_proxy_async();  // generic-proxy accesses to rec are ordered before the next bulk load
    g.sync();
  }
  if (!legmode) {
    if (g.tid == 0) tma_store_wait_read();
    g.sync();
    initial_solve(cx, p, b);
  }
  // non-finite check on the running value function (cheap proxy for the whole sweep)
  for (int e = g.tid; e < n * n; e += NT)
    if (!isfinite(cx.P[e])) cx.flags |= GARB200_FLAG_NONFINITE;
  if (cx.flags) atomicOr(reinterpret_cast<int *>(p.flags) + b, cx.flags);
  if (g.tid == 0) tma_store_wait_all();
}

template <int NT, int TM, int TN>
static int launch_backward_t(const DevParams &p, cudaStream_t st) {
  const SmemPlan pl = make_plan(p);
  const size_t smem = (size_t)pl.total * sizeof(double);
  if (smem > 227u * 1024u) return -1;
  auto kern = backward_kernel<NT, TM, TN>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return -2;
  const int nlegs = p.num_legs >= 2 ? p.num_legs : 1;
  kern<<<p.batch * nlegs, NT, smem, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

int launch_backward(const DevParams &p, cudaStream_t st) {
  if (p.nx <= 16) return launch_backward_t<32, 3, 2>(p, st);
  if (p.nx <= 48) return launch_backward_t<128, 4, 3>(p, st);
  return launch_backward_t<256, 4, 4>(p, st);
}

size_t backward_smem_bytes(const DevParams &p) { return (size_t)make_plan(p).total * sizeof(double); }

} // namespace garb200
