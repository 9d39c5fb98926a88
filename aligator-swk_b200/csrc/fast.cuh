// fast.cuh — compile-time-dimension building blocks of the specialised sweep (sm_100a).
//
// Everything here is templated on the problem dimensions so that loops unroll, strides fold
// into immediates and no integer division survives.  Dense contractions run on the FP64
// tensor pipe (DMMA, mma.sync.m8n8k4.f64 — tcgen05 has no FP64 kind); factorizations are
// hand-scheduled on shared memory with one barrier per pivot.
//
//   mma_gemm     every A'PA / B'PB / K'QuuK-class product of stageKernelSolve
//                (riccati-kernel.hxx:175-199, 238-250) incl. the GEMV that rides along as an
//                extra column
//   ldlt_solve   Eigen::LLT(I + mu*Ptilde).solveInPlace      riccati-kernel.hxx:185-189
//   lu_inverse_warp  Eigen::PartialPivLU(E).solve(I)         riccati-kernel.hxx:172-174
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>
#include "device_linalg.cuh"

namespace garb200 {

__device__ __forceinline__ void dmma_884(double &c0, double &c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c0), "+d"(c1)
      : "d"(a), "d"(b));
}

// ------------------------------------------------------------------------------------------
// C(i,j) = cin(i,j) + sum_k a(i,k) b(k,j),  M x N x K, distributed over NW warps.
// a / b / cin / store are functors (lambdas): a(i,k), b(k,j), cin(i,j) -> double,
// store(i,j,v).  Indices handed to a/b/cin are always clamped into range; padded rows /
// columns of the 8x8 tiles compute garbage that is never stored.  K is zero-filled to a
// multiple of 4.
// No barrier inside: the caller synchronises before (operands ready) and after.
// ------------------------------------------------------------------------------------------
template <int M, int N, int K, int NW, class FA, class FB, class FC, class FS>
__device__ __forceinline__ void mma_gemm(int warp, int lane, FA a, FB b, FC cin, FS store) {
  if constexpr (M <= 0 || N <= 0) return;
  constexpr int TM = (M + 7) / 8, TN = (N + 7) / 8, NTILE = TM * TN;
  constexpr int NSLOT = (NTILE + NW - 1) / NW;      // tiles per warp, ALL kept in flight:
  constexpr int KS = (K + 3) / 4;                   // NSLOT independent DMMA accumulation chains
  const int g = lane >> 2, q = lane & 3;
  // Strip mode: when the row tiles (or else the column tiles) divide evenly over the warps,
  // a warp owns whole 8-row strips (8-column strips) of C, so one A (B) fragment feeds every
  // tile of the strip: (SR + SC) shared-memory loads per SR * SC DMMAs instead of 2 per DMMA.
  constexpr bool ROWS = (TM % NW == 0);
  constexpr bool COLS = !ROWS && (TN % NW == 0);
  if constexpr (ROWS || COLS) {
    constexpr int SR = ROWS ? TM / NW : TM, SC = ROWS ? TN : TN / NW;
    double acc[SR][SC][2];
    int ia[SR], jb[SC];
#pragma unroll
    for (int r = 0; r < SR; r++) ia[r] = min((ROWS ? warp + NW * r : r) * 8 + g, M - 1);
#pragma unroll
    for (int u = 0; u < SC; u++) jb[u] = min((ROWS ? u : warp + NW * u) * 8 + g, N - 1);
#pragma unroll
    for (int r = 0; r < SR; r++)
#pragma unroll
      for (int u = 0; u < SC; u++) {
        const int jc = (ROWS ? u : warp + NW * u) * 8 + 2 * q;
        acc[r][u][0] = cin(ia[r], min(jc, N - 1));
        acc[r][u][1] = cin(ia[r], min(jc + 1, N - 1));
      }
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      const int k = ks * 4 + q;
      double av[SR], bv[SC];
#pragma unroll
      for (int r = 0; r < SR; r++) {
        if constexpr (K % 4 == 0) av[r] = a(ia[r], k);
        else av[r] = (k < K) ? a(ia[r], min(k, K - 1)) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < SC; u++) {
        if constexpr (K % 4 == 0) bv[u] = b(k, jb[u]);
        else bv[u] = (k < K) ? b(min(k, K - 1), jb[u]) : 0.0;
      }
#pragma unroll
      for (int r = 0; r < SR; r++)
#pragma unroll
        for (int u = 0; u < SC; u++) dmma_884(acc[r][u][0], acc[r][u][1], av[r], bv[u]);
    }
#pragma unroll
    for (int r = 0; r < SR; r++)
#pragma unroll
      for (int u = 0; u < SC; u++) {
        const int ic = (ROWS ? warp + NW * r : r) * 8 + g, jc = (ROWS ? u : warp + NW * u) * 8 + 2 * q;
        if (ic < M) {
          if (jc < N) store(ic, jc, acc[r][u][0]);
          if (jc + 1 < N) store(ic, jc + 1, acc[r][u][1]);
        }
      }
  } else {
  double acc[NSLOT][2];
  int ia[NSLOT], jb[NSLOT];
#pragma unroll
  for (int u = 0; u < NSLOT; u++) {
    const int t = min(warp + NW * u, NTILE - 1);
    const int ti = t % TM, tj = t / TM;
    ia[u] = min(ti * 8 + g, M - 1);
    jb[u] = min(tj * 8 + g, N - 1);
    const int jc = tj * 8 + 2 * q;
    acc[u][0] = cin(ia[u], min(jc, N - 1));
    acc[u][1] = cin(ia[u], min(jc + 1, N - 1));
  }
#pragma unroll
  for (int ks = 0; ks < KS; ks++) {
    const int k = ks * 4 + q;
    double av[NSLOT], bv[NSLOT];
#pragma unroll
    for (int u = 0; u < NSLOT; u++) {
      if constexpr (K % 4 == 0) { av[u] = a(ia[u], k); bv[u] = b(k, jb[u]); }
      else {
        av[u] = (k < K) ? a(ia[u], min(k, K - 1)) : 0.0;
        bv[u] = (k < K) ? b(min(k, K - 1), jb[u]) : 0.0;
      }
    }
#pragma unroll
    for (int u = 0; u < NSLOT; u++) dmma_884(acc[u][0], acc[u][1], av[u], bv[u]);
  }
#pragma unroll
  for (int u = 0; u < NSLOT; u++) {
    const int t = warp + NW * u;
    if (t < NTILE) {
      const int ti = t % TM, tj = t / TM;
      const int ic = ti * 8 + g, jc = tj * 8 + 2 * q;
      if (ic < M) {
        if (jc < N) store(ic, jc, acc[u][0]);
        if (jc + 1 < N) store(ic, jc + 1, acc[u][1]);
      }
    }
  }
  }
}

// ------------------------------------------------------------------------------------------
// Column-strip GEMM that may run IN PLACE on its b operand:  C(:, J) <- cin(:, J) + a b(:, J).
// A warp owns whole 8-column strips of C (strip J -> warp J % NW), computes all row tiles of a strip
// before storing any of it, and is the only warp that reads those columns of b — so `store` may
// overwrite b.  a and cin must not alias the output.  One b fragment feeds TM DMMAs.
// ------------------------------------------------------------------------------------------
template <int M, int N, int K, int NW, class FA, class FB, class FC, class FS>
__device__ __forceinline__ void mma_gemm_colstrips(int warp, int lane, FA a, FB b, FC cin, FS store) {
  constexpr int TM = (M + 7) / 8, TN = (N + 7) / 8, KS = (K + 3) / 4;
  const int g = lane >> 2, q = lane & 3;
  int ia[TM];
#pragma unroll
  for (int r = 0; r < TM; r++) ia[r] = min(r * 8 + g, M - 1);
#pragma unroll 1
  for (int tj = warp; tj < TN; tj += NW) {
    const int jb = min(tj * 8 + g, N - 1), jc = tj * 8 + 2 * q;
    double acc[TM][2];
#pragma unroll
    for (int r = 0; r < TM; r++) {
      acc[r][0] = cin(ia[r], min(jc, N - 1));
      acc[r][1] = cin(ia[r], min(jc + 1, N - 1));
    }
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      const int k = ks * 4 + q;
      const bool kv = (K % 4 == 0) || (k < K);
      const int kc = min(k, K - 1);
      const double bv = kv ? b(kc, jb) : 0.0;
#pragma unroll
      for (int r = 0; r < TM; r++) dmma_884(acc[r][0], acc[r][1], kv ? a(ia[r], kc) : 0.0, bv);
    }
    __syncwarp();                                  // every lane has read its b fragments of this strip
#pragma unroll
    for (int r = 0; r < TM; r++) {
      const int ic = r * 8 + g;
      if (ic < M) {
        if (jc < N) store(ic, jc, acc[r][0]);
        if (jc + 1 < N) store(ic, jc + 1, acc[r][1]);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Row-strip GEMM for operands that do not all fit in shared memory (nx = 76): a warp owns whole
// 8-row strips of C (strip s -> warp s % NW, or the mirrored assignment with REV so that two
// back-to-back uneven GEMMs load the warps evenly).  The strip-side operand a(i, k) may live in
// GLOBAL memory: with PRE its K/4 fragments of the strip are fetched into registers before the
// first DMMA issues (one L2 round trip per strip instead of one per k-step), and every element of
// it is read exactly once per GEMM.  b(k, j) is read by every warp and must be in shared memory.
// ------------------------------------------------------------------------------------------
template <int M, int N, int K, int NW, bool PRE, bool REV, class FA, class FB, class FC, class FS>
__device__ __forceinline__ void strip_gemm(int warp, int lane, FA a, FB b, FC cin, FS store) {
  constexpr int TM = (M + 7) / 8, TN = (N + 7) / 8, KS = (K + 3) / 4;
  const int g = lane >> 2, q = lane & 3;
  const int w0 = REV ? (NW - 1 - warp) : warp;
  int jb[TN];
#pragma unroll
  for (int u = 0; u < TN; u++) jb[u] = min(u * 8 + g, N - 1);
#pragma unroll 1
  for (int sidx = w0; sidx < TM; sidx += NW) {
    const int ic = sidx * 8 + g, ia = min(ic, M - 1);
    double av[PRE ? KS : 1];
    if constexpr (PRE) {
#pragma unroll
      for (int ks = 0; ks < KS; ks++) {
        const int k = ks * 4 + q;
        av[ks] = (K % 4 == 0 || k < K) ? a(ia, min(k, K - 1)) : 0.0;
      }
    }
    double acc[TN][2];
#pragma unroll
    for (int u = 0; u < TN; u++) {
      const int jc = u * 8 + 2 * q;
      acc[u][0] = cin(ia, min(jc, N - 1));
      acc[u][1] = cin(ia, min(jc + 1, N - 1));
    }
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      const int k = ks * 4 + q;
      const bool kv = (K % 4 == 0) || (k < K);
      const int kc = min(k, K - 1);
      double a1;
      if constexpr (PRE) a1 = av[ks]; else a1 = kv ? a(ia, kc) : 0.0;
      double bv[TN];
#pragma unroll
      for (int u = 0; u < TN; u++) bv[u] = kv ? b(kc, jb[u]) : 0.0;
#pragma unroll
      for (int u = 0; u < TN; u++) dmma_884(acc[u][0], acc[u][1], a1, bv[u]);
    }
    if (ic < M) {
#pragma unroll
      for (int u = 0; u < TN; u++) {
        const int jc = u * 8 + 2 * q;
        if (jc < N) store(ic, jc, acc[u][0]);
        if (jc + 1 < N) store(ic, jc + 1, acc[u][1]);
      }
    }
  }
}

// Symmetric product: only the 8x8 tiles on or below the diagonal of the N x N result are
// computed (TM (TM + 1) / 2 instead of TM^2 tiles), dealt round-robin to the warps.  `store` still
// sees every element of a diagonal tile and filters i >= j itself.  (Skipping the upper tiles of
// Qhat and Pmat the same way was measured and lost: their mirrored stores cost more than the
// DMMAs saved.)
template <int N, int K, int NW, class FA, class FB, class FC, class FS>
__device__ __forceinline__ void mma_gemm_lower(int warp, int lane, FA a, FB b, FC cin, FS store) {
  constexpr int TM = (N + 7) / 8, NTILE = TM * (TM + 1) / 2;
  constexpr int NSLOT = (NTILE + NW - 1) / NW, KS = (K + 3) / 4;
  static_assert(K % 4 == 0 && TM <= 8, "mma_gemm_lower: K % 4 == 0, N <= 64");
  const int g = lane >> 2, q = lane & 3;
  double acc[NSLOT][2];
  int ia[NSLOT], jb[NSLOT], ti_[NSLOT], tj_[NSLOT];
#pragma unroll
  for (int u = 0; u < NSLOT; u++) {
    const int t = min(warp + NW * u, NTILE - 1);
    int ti = 0;
#pragma unroll
    for (int r = 1; r < TM; r++) if (t >= r * (r + 1) / 2) ti = r;      // row of the packed lower triangle
    const int tj = t - ti * (ti + 1) / 2;
    ti_[u] = ti; tj_[u] = tj;
    ia[u] = min(ti * 8 + g, N - 1);
    jb[u] = min(tj * 8 + g, N - 1);
    const int jc = tj * 8 + 2 * q;
    acc[u][0] = cin(ia[u], min(jc, N - 1));
    acc[u][1] = cin(ia[u], min(jc + 1, N - 1));
  }
#pragma unroll
  for (int ks = 0; ks < KS; ks++) {
    const int k = ks * 4 + q;
    double av[NSLOT], bv[NSLOT];
#pragma unroll
    for (int u = 0; u < NSLOT; u++) { av[u] = a(ia[u], k); bv[u] = b(k, jb[u]); }
#pragma unroll
    for (int u = 0; u < NSLOT; u++) dmma_884(acc[u][0], acc[u][1], av[u], bv[u]);
  }
#pragma unroll
  for (int u = 0; u < NSLOT; u++) {
    if (warp + NW * u < NTILE) {
      const int ic = ti_[u] * 8 + g, jc = tj_[u] * 8 + 2 * q;
      if (ic < N) {
        if (jc < N) store(ic, jc, acc[u][0]);
        if (jc + 1 < N) store(ic, jc + 1, acc[u][1]);
      }
    }
  }
}

// y(i) = yin(i) + alpha * sum_j A(i,j) x(j): 4 threads per row + 2 shuffles; A col-major.
// Threads beyond 4*M idle.  NT must be a multiple of 32.  fin / fout are functors.
template <int M, int K, int NT, class FA, class FX, class FI, class FO>
__device__ __forceinline__ void gemv4(int tid, FA a, FX x, FI yin, FO yout) {
  const int q = tid & 3;
#pragma unroll 1
  for (int r0 = 0; r0 < M; r0 += NT / 4) {
    const int r = r0 + (tid >> 2);
    const int rc = min(r, M - 1);
    double acc = 0.0;
#pragma unroll
    for (int j0 = 0; j0 < K; j0 += 4) {
      const int j = j0 + q;
      if (K % 4 == 0 || j < K) acc = fma(a(rc, min(j, K - 1)), x(min(j, K - 1)), acc);
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (r < M && q == 0) yout(r, yin(r) + acc);
  }
}


// A(j,i) = A(i,j), i > j (col-major, ld = N): selfadjointView<Lower>
template <int N, int NT> __device__ __forceinline__ void mirror_lower_ct(int tid, double *A) {
  for (int e = tid; e < N * N; e += NT) {
    const int i = e % N, j = e / N;
    if (i > j) A[j + i * N] = A[i + j * N];
  }
}

// ------------------------------------------------------------------------------------------
// Runtime-extent DMMA update with ONE k-step (K = 4):  C(i,j) = cin(i,j) + sum_{k<4} a(i,k) b(k,j)
// for i < M, j < N; `lower`: only the 8x8 tiles on or below the diagonal (M == N).
// ------------------------------------------------------------------------------------------
template <int NW, int TNMAX, class FA, class FB, class FC, class FS>
__device__ __forceinline__ void mma_k4(int warp, int lane, int M, int N, bool lower, FA a, FB b, FC cin,
                                       FS store) {
  if (M <= 0 || N <= 0) return;
  const int TM = (M + 7) >> 3, TN = (N + 7) >> 3;
  const int g = lane >> 2, q = lane & 3;
  // one warp per 8-row strip: the A fragment is loaded once, every operand of the strip is
  // fetched before the first DMMA issues (loads, tensor ops and stores overlap across tiles)
  for (int ti = warp; ti < TM; ti += NW) {
    const int ic = ti * 8 + g, ia = min(ic, M - 1);
    const double av = a(ia, q);
    double c0[TNMAX], c1[TNMAX], bv[TNMAX];
#pragma unroll
    for (int u = 0; u < TNMAX; u++) {
      const bool on = u < TN && (!lower || u <= ti);
      if (on) {
        const int jb = min(u * 8 + g, N - 1), jc = u * 8 + 2 * q;
        bv[u] = b(q, jb);
        c0[u] = cin(ia, min(jc, N - 1));
        c1[u] = cin(ia, min(jc + 1, N - 1));
      }
    }
#pragma unroll
    for (int u = 0; u < TNMAX; u++) {
      const bool on = u < TN && (!lower || u <= ti);
      if (on) dmma_884(c0[u], c1[u], av, bv[u]);
    }
#pragma unroll
    for (int u = 0; u < TNMAX; u++) {
      const bool on = u < TN && (!lower || u <= ti);
      const int jc = u * 8 + 2 * q;
      if (on && ic < M) {
        if (jc < N) store(ic, jc, c0[u]);
        if (jc + 1 < N) store(ic, jc + 1, c1[u]);
      }
    }
  }
}

/// Right-hand-side accessors for ldlt_solve_blocked: element (i, j) of the N x NR block.
template <int LD> struct RhsColMajor {
  double *p;
  __device__ __forceinline__ double get(int i, int j) const { return p[i + j * LD]; }
  __device__ __forceinline__ void set(int i, int j, double v) const { p[i + j * LD] = v; }
};
/// The same with the theta block and padding: column j > n is column j - n - 1 of the row-major
/// fth rows; rows >= S (padding of the blocked factorisation) read 0 and drop writes.
template <int NXc, int H, int S> struct RhsGainsTh {
  double *ff, *fb, *fth;
  __device__ __forceinline__ double get(int i, int j) const {
    if (i >= S) return 0.0;
    return j < NXc ? fb[i * NXc + j] : (j == NXc ? ff[i] : fth[i * H + (j - NXc - 1)]);
  }
  __device__ __forceinline__ void set(int i, int j, double v) const {
    if (i >= S) return;
    if (j < NXc) fb[i * NXc + j] = v; else if (j == NXc) ff[i] = v; else fth[i * H + (j - NXc - 1)] = v;
  }
};

// ------------------------------------------------------------------------------------------
// Solve S X = R in place by a BLOCKED square-root-free LDL^T (N % 4 == 0, lower triangle of the
// col-major S read and overwritten by L; R: N x NR right-hand sides through an accessor).
// Four pivots per block step: (A) every thread factors the 4x4 diagonal block redundantly in
// registers, threads < N-r0-4 turn their row of the panel into L and keep the unscaled row in
// Wp, threads CO.. forward-substitute one right-hand-side column each; (B) the trailing matrix and
// the right-hand sides below the block are updated with K = 4 DMMA tiles (mma_k4).  Backward
// substitution mirrors it (C: 4 rows per column in registers, D: DMMA update of the rows above).
// The pivot signs are checked against (+)^NPOS (-)^(N-NPOS): returns false when the matrix does
// not have that inertia (SPD: NPOS = N; quasi-definite KKT block: NPOS = nu).
// Scratch: Wp 4 (N - 4) doubles, dinv N doubles.  Needs NT >= round32(N - 4) + NR threads.
// ------------------------------------------------------------------------------------------
template <int N, int NR, int NT, int NPOS = N, int GRP = 1, class RH>
__device__ __forceinline__ bool ldlt_solve_blocked(int tid, double *S, RH R, double *Wp, double *dinv) {
  static_assert(N % 4 == 0, "blocked LDL^T needs N % 4 == 0");
  constexpr int NW = NT / 32;
  constexpr int CO_ = ((N - 4 + 31) / 32) * 32;
  constexpr int CO = (CO_ + NR <= NT) ? CO_ : (N - 4 > 0 ? N - 4 : 0);
  static_assert(CO + NR <= NT, "not enough threads for rows + rhs columns");
  const int warp = tid >> 5, lane = tid & 31;
  bool ok = true;
#pragma unroll 1
  for (int r0 = 0; r0 < N; r0 += 4) {
    // ---- phase A
    const double *Dg = S + r0 + r0 * N;
    const double a00 = Dg[0], a10 = Dg[1], a20 = Dg[2], a30 = Dg[3];
    const double a11 = Dg[1 + N], a21 = Dg[2 + N], a31 = Dg[3 + N];
    const double a22 = Dg[2 + 2 * N], a32 = Dg[3 + 2 * N], a33 = Dg[3 + 3 * N];
    // The reciprocal pivots come in PAIRS (2x2 leading minors): 1/d1 = a00 / (a00 a11 - a10^2)
    // does not wait for 1/d0, which shortens the dependent FP64 chain of the block from ~30 to
    // ~20 operations.  Dependent FP64 latency is what this phase costs: the other CTAs of the SM
    // keep the FP64 pipe full of DMMAs, so every dependent DFMA waits ~100 cycles (measured).
    const double i0 = __drcp_rn(a00);
    const double det01 = fma(a00, a11, -a10 * a10);
    const double i1 = a00 * __drcp_rn(det01);
    const double l10 = a10 * i0, l20 = a20 * i0, l30 = a30 * i0;
    const double d1 = det01 * i0;
    const double w21 = fma(-l20, a10, a21), w31 = fma(-l30, a10, a31);     // unscaled
    const double l21 = w21 * i1, l31 = w31 * i1;
    const double d2 = fma(-l21, w21, fma(-l20, a20, a22));
    const double w32 = fma(-l31, w21, fma(-l30, a20, a32));
    const double s33 = fma(-l31, w31, fma(-l30, a30, a33));
    const double i2 = __drcp_rn(d2);
    const double det23 = fma(d2, s33, -w32 * w32);
    const double i3 = d2 * __drcp_rn(det23);
    const double l32 = w32 * i2;
    const double d3 = det23 * i2;
    // inertia check: the first NPOS pivots must be > 0, the others < 0 (quasi-definite matrix)
    {
      const double s0 = (r0 + 0 < NPOS) ? a00 : -a00, s1 = (r0 + 1 < NPOS) ? d1 : -d1;
      const double s2 = (r0 + 2 < NPOS) ? d2 : -d2, s3 = (r0 + 3 < NPOS) ? d3 : -d3;
      if (!(s0 > 0.0) || !(s1 > 0.0) || !(s2 > 0.0) || !(s3 > 0.0)) ok = false;
    }
    const int nbelow = N - r0 - 4;
    if (tid < nbelow) {
      const int i = r0 + 4 + tid;
      double *row = S + i + r0 * N;
      const double w0 = row[0];
      const double l0 = w0 * i0;
      const double w1 = fma(-l0, a10, row[N]);
      const double l1 = w1 * i1;
      const double w2 = fma(-l1, w21, fma(-l0, a20, row[2 * N]));
      const double l2 = w2 * i2;
      const double w3 = fma(-l2, w32, fma(-l1, w31, fma(-l0, a30, row[3 * N])));
      const double l3 = w3 * i3;
      row[0] = l0; row[N] = l1; row[2 * N] = l2; row[3 * N] = l3;
      double *wp = Wp + 4 * (i - r0 - 4);
      wp[0] = w0; wp[1] = w1; wp[2] = w2; wp[3] = w3;
    } else if (tid >= CO && tid < CO + NR) {
      const int j = tid - CO;
      const double y0 = R.get(r0, j);
      const double y1 = fma(-l10, y0, R.get(r0 + 1, j));
      const double y2 = fma(-l21, y1, fma(-l20, y0, R.get(r0 + 2, j)));
      const double y3 = fma(-l32, y2, fma(-l31, y1, fma(-l30, y0, R.get(r0 + 3, j))));
      R.set(r0 + 1, j, y1); R.set(r0 + 2, j, y2); R.set(r0 + 3, j, y3);
    }
    cta_sync<NT, GRP>();
    if (tid == NT - 1) {   // after the barrier: every thread has read the unscaled diagonal block
      double *Dw = S + r0 + r0 * N;
      Dw[1] = l10; Dw[2] = l20; Dw[3] = l30; Dw[2 + N] = l21; Dw[3 + N] = l31; Dw[3 + 2 * N] = l32;
      dinv[r0] = i0; dinv[r0 + 1] = i1; dinv[r0 + 2] = i2; dinv[r0 + 3] = i3;
    }
    // ---- phase B (rows/cols below the block, origin r1)
    if (nbelow > 0) {
      const int r1 = r0 + 4;
      const double *Lp = S + r1 + r0 * N;          // L panel: (i, k) at Lp[i + k*N]
      mma_k4<NW, (NR + 7) / 8>(warp, lane, nbelow, NR, false,
                 [&](int i, int k) { return -Lp[i + k * N]; },
                 [&](int k, int j) { return R.get(r0 + k, j); },
                 [&](int i, int j) { return R.get(r1 + i, j); },
                 [&](int i, int j, double v) { R.set(r1 + i, j, v); });
      double *S1 = S + r1 + r1 * N;
      mma_k4<NW, ((N + 3) / 8 < 1 ? 1 : (N + 3) / 8)>(warp, lane, nbelow, nbelow, true,
                 [&](int i, int k) { return -Lp[i + k * N]; },
                 [&](int k, int j) { return Wp[4 * j + k]; },
                 [&](int i, int j) { return S1[i + j * N]; },
                 [&](int i, int j, double v) { S1[i + j * N] = v; });
      cta_sync<NT, GRP>();
    }
  }
  cta_sync<NT, GRP>();
  // z = D^{-1} y
  for (int e = tid; e < N * NR; e += NT) { const int i = e % N, j = e / N; R.set(i, j, R.get(i, j) * dinv[i]); }
  cta_sync<NT, GRP>();
#pragma unroll 1
  for (int r0 = N - 4; r0 >= 0; r0 -= 4) {
    // ---- phase C: x rows r0..r0+3 of every column
    if (tid < NR) {
      const double *Dg = S + r0 + r0 * N;
      const double l10 = Dg[1], l20 = Dg[2], l30 = Dg[3], l21 = Dg[2 + N], l31 = Dg[3 + N], l32 = Dg[3 + 2 * N];
      const double x3 = R.get(r0 + 3, tid);
      const double x2 = fma(-l32, x3, R.get(r0 + 2, tid));
      const double x1 = fma(-l31, x3, fma(-l21, x2, R.get(r0 + 1, tid)));
      const double x0 = fma(-l30, x3, fma(-l20, x2, fma(-l10, x1, R.get(r0, tid))));
      R.set(r0 + 2, tid, x2); R.set(r0 + 1, tid, x1); R.set(r0, tid, x0);
    }
    if (r0 == 0) break;
    cta_sync<NT, GRP>();
    // ---- phase D: rows above: R(i, j) -= sum_k L(r0+k, i) x(r0+k, j)
    const double *Lr = S + r0;                     // (r0+k, i) at Lr[k + i*N]
    mma_k4<NW, (NR + 7) / 8>(warp, lane, r0, NR, false,
               [&](int i, int k) { return -Lr[k + i * N]; },
               [&](int k, int j) { return R.get(r0 + k, j); },
               [&](int i, int j) { return R.get(i, j); },
               [&](int i, int j, double v) { R.set(i, j, v); });
    cta_sync<NT, GRP>();
  }
  cta_sync<NT, GRP>();
  return ok;
}

// ------------------------------------------------------------------------------------------
// The same blocked LDL^T split into its two halves so that the FACTORIZATION (a dependent FP64
// chain: ~30 levels per 4 pivots, no parallelism to speak of) runs on ONE warp beside a GEMM of
// the other warps, and the SUBSTITUTIONS run one right-hand-side column per thread with the
// column in registers (depth 2 N FMAs, no barrier, L read as shared-memory broadcasts).
//
// ldlt_factor_warp: S (N x N col-major, lower triangle) -> unit-lower L in place, dinv[N].
//   One warp, __syncwarp only.  Wp: 4 (N - 4) doubles of scratch.
//   Returns false when the pivots do not have the inertia (+)^NPOS (-)^(N-NPOS), or when a pivot
//   of the leading (positive) block is numerically zero relative to the largest diagonal entry
//   (|d_k| <= 1e-13 max_i |S_ii|: a rank-deficient Rhat whose controls are fixed by the constraint
//   rows only — the unpivoted order would divide by rounding noise; the caller then re-factorises
//   with the reference's Bunch-Kaufman pivoting, bunchkaufman.hpp:20-167).
// ------------------------------------------------------------------------------------------
template <int N, int NPOS = N>
__device__ __forceinline__ bool ldlt_factor_warp(int lane, double *S, double *Wp, double *dinv) {
  static_assert(N % 4 == 0 && N - 4 <= 32, "in-warp blocked LDL^T: N % 4 == 0, N <= 36");
  bool ok = true;
  double amax = 0.0;
#pragma unroll
  for (int i = 0; i < NPOS; i++) amax = fmax(amax, fabs(S[i + i * N]));
  const double tiny = 1e-13 * amax;
#pragma unroll 1
  for (int r0 = 0; r0 < N; r0 += 4) {
    const double *Dg = S + r0 + r0 * N;
    const double a00 = Dg[0], a10 = Dg[1], a20 = Dg[2], a30 = Dg[3];
    const double a11 = Dg[1 + N], a21 = Dg[2 + N], a31 = Dg[3 + N];
    const double a22 = Dg[2 + 2 * N], a32 = Dg[3 + 2 * N], a33 = Dg[3 + 3 * N];
    // reciprocal pivots in pairs (2x2 leading minors), see ldlt_solve_blocked
    const double i0 = __drcp_rn(a00);
    const double det01 = fma(a00, a11, -a10 * a10);
    const double i1 = a00 * __drcp_rn(det01);
    const double l10 = a10 * i0, l20 = a20 * i0, l30 = a30 * i0;
    const double d1 = det01 * i0;
    const double w21 = fma(-l20, a10, a21), w31 = fma(-l30, a10, a31);
    const double l21 = w21 * i1, l31 = w31 * i1;
    const double d2 = fma(-l21, w21, fma(-l20, a20, a22));
    const double w32 = fma(-l31, w21, fma(-l30, a20, a32));
    const double s33 = fma(-l31, w31, fma(-l30, a30, a33));
    const double i2 = __drcp_rn(d2);
    const double det23 = fma(d2, s33, -w32 * w32);
    const double i3 = d2 * __drcp_rn(det23);
    const double l32 = w32 * i2;
    const double d3 = det23 * i2;
    {
      const double dd[4] = {a00, d1, d2, d3};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const bool pos = r0 + k < NPOS;
        const double sk = pos ? dd[k] : -dd[k];
        if (!(sk > 0.0) || (pos && !(sk > tiny))) ok = false;
      }
    }
    const int nbelow = N - r0 - 4;
    if (lane < nbelow) {
      const int i = r0 + 4 + lane;
      double *row = S + i + r0 * N;
      const double w0 = row[0];
      const double l0 = w0 * i0;
      const double w1 = fma(-l0, a10, row[N]);
      const double l1 = w1 * i1;
      const double w2 = fma(-l1, w21, fma(-l0, a20, row[2 * N]));
      const double l2 = w2 * i2;
      const double w3 = fma(-l2, w32, fma(-l1, w31, fma(-l0, a30, row[3 * N])));
      const double l3 = w3 * i3;
      row[0] = l0; row[N] = l1; row[2 * N] = l2; row[3 * N] = l3;
      double *wp = Wp + 4 * lane;
      wp[0] = w0; wp[1] = w1; wp[2] = w2; wp[3] = w3;
    }
    __syncwarp();
    if (lane == 31) {   // every lane has read the unscaled diagonal block
      double *Dw = S + r0 + r0 * N;
      Dw[1] = l10; Dw[2] = l20; Dw[3] = l30; Dw[2 + N] = l21; Dw[3 + N] = l31; Dw[3 + 2 * N] = l32;
      dinv[r0] = i0; dinv[r0 + 1] = i1; dinv[r0 + 2] = i2; dinv[r0 + 3] = i3;
    }
    if (nbelow > 0) {
      const int r1 = r0 + 4;
      const double *Lp = S + r1 + r0 * N;
      double *S1 = S + r1 + r1 * N;
      mma_k4<1, ((N + 3) / 8 < 1 ? 1 : (N + 3) / 8)>(0, lane, nbelow, nbelow, true,
                 [&](int i, int k) { return -Lp[i + k * N]; },
                 [&](int k, int j) { return Wp[4 * j + k]; },
                 [&](int i, int j) { return S1[i + j * N]; },
                 [&](int i, int j, double v) { S1[i + j * N] = v; });
    }
    __syncwarp();
  }
  return ok;
}

// ldlt_solve_cols: X <- (L D L^T)^{-1} X for the NR columns of the accessor R, one column per
// thread (threads tid < NR), the column held in registers.  Column-oriented (axpy) substitutions:
// dependent depth 2 N, every L entry is one shared-memory broadcast.  No barrier inside; the
// caller synchronises before (factor + right-hand sides complete) and after.
template <int N, int NR, class RH>
__device__ __forceinline__ void ldlt_solve_cols(int tid, const double *S, const double *dinv, RH R) {
  if (tid >= NR) return;
  // (volatile: the L entries are consumed in program order — left free, the scheduler of the fully
  //  unrolled N = 32 form hoists all 496 loads and spills 4 KB per thread)
  const volatile double *L = S;
  double x[N];
#pragma unroll
  for (int i = 0; i < N; i++) x[i] = R.get(i, tid);
#pragma unroll
  for (int k = 0; k < N - 1; k++) {
#pragma unroll
    for (int i = k + 1; i < N; i++) x[i] = fma(-L[i + k * N], x[k], x[i]);
  }
#pragma unroll
  for (int i = 0; i < N; i++) x[i] *= dinv[i];
#pragma unroll
  for (int k = N - 1; k > 0; k--) {
#pragma unroll
    for (int i = 0; i < k; i++) x[i] = fma(-L[k + i * N], x[k], x[i]);
  }
#pragma unroll
  for (int i = 0; i < N; i++) R.set(i, tid, x[i]);
}

// ldlt_solve_lanes: the same substitutions with every right-hand-side column spread over LPC
// lanes of a warp (LPC = 4, 2 or 1: the largest with LPC * NR <= NT).  Lane q of a column's group
// owns rows q, q + LPC, ...; at step k the owner broadcasts x_k with one shuffle and every lane
// updates its own rows: still 2 N dependent steps, but each is one shuffle + ONE predicated FMA
// deep instead of N - k FMAs, and LPC * NR threads (all warps of the CTA) work instead of NR.
// All lanes of a participating warp execute the shuffles (lanes past the last column run a clamped
// duplicate and store nothing).  No barrier inside.
template <int N, int NR, int NT, class RH>
__device__ __forceinline__ void ldlt_solve_lanes(int tid, const double *S, const double *dinv, RH R) {
  constexpr int LPC = (4 * NR <= NT) ? 4 : (2 * NR <= NT) ? 2 : 1;
  if constexpr (LPC == 1) { ldlt_solve_cols<N, NR>(tid, S, dinv, R); return; }
  constexpr int RP = (N + LPC - 1) / LPC;          // rows per lane
  if ((tid & ~31) >= LPC * NR) return;             // whole warps past the last column
  const int lane = tid & 31, q = lane & (LPC - 1), gbase = lane & ~(LPC - 1);
  const bool live = tid < LPC * NR;
  const int col = live ? tid / LPC : NR - 1;
  using LT = typename std::conditional<(N > 16), const volatile double, const double>::type;
  LT *L = S;                                       // (large N: keep the loads in program order, see ldlt_solve_cols)
  double x[RP];
#pragma unroll
  for (int r = 0; r < RP; r++) { const int i = r * LPC + q; x[r] = i < N ? R.get(i, col) : 0.0; }
#pragma unroll
  for (int k = 0; k < N - 1; k++) {
    const double xk = __shfl_sync(0xffffffffu, x[k / LPC], gbase + (k % LPC));
#pragma unroll
    for (int r = k / LPC; r < RP; r++) {
      const int i = r * LPC + q;
      if (i > k && i < N) x[r] = fma(-L[i + k * N], xk, x[r]);
    }
  }
#pragma unroll
  for (int r = 0; r < RP; r++) { const int i = r * LPC + q; if (i < N) x[r] *= dinv[i]; }
#pragma unroll
  for (int k = N - 1; k > 0; k--) {
    const double xk = __shfl_sync(0xffffffffu, x[k / LPC], gbase + (k % LPC));
#pragma unroll
    for (int r = 0; r <= k / LPC; r++) {
      const int i = r * LPC + q;
      if (i < k) x[r] = fma(-L[k + i * N], xk, x[r]);
    }
  }
  if (live) {
#pragma unroll
    for (int r = 0; r < RP; r++) { const int i = r * LPC + q; if (i < N) R.set(i, col, x[r]); }
  }
}

// ------------------------------------------------------------------------------------------
// ldlt_solve_dmma: the same two substitutions BLOCKED on the tensor pipe.  A warp owns whole 8-column
// tiles of the right-hand sides and keeps all N rows of its tile in registers as m8n8k4 accumulator
// fragments (lane (g, q): rows g + 8 mt, columns 2q, 2q + 1).  Per 4-row block: (a) the 4 x 4 unit
// triangular block is substituted in registers (the four rows sit in four lanes of equal q: one
// shuffle + one FMA per dependent step), (b) the finished block becomes the B fragment of ONE rank-4
// DMMA per row tile still to be updated (C -> B layout by one shuffle pair + select), the A fragments
// (-L) come straight from the factor in shared memory.  2 N / 4 dependent block steps of ~10
// instructions instead of 2 N shuffle + FMA steps, and no barrier: column tiles are independent.
// S: unit lower factor (col-major, leading dimension N, N % 4 == 0), dinv: 1 / d.
// ------------------------------------------------------------------------------------------
template <int N, int NR, int NT, class RH>
__device__ __forceinline__ void ldlt_solve_dmma(int tid, const double *S, const double *dinv, RH R) {
  static_assert(N % 4 == 0, "ldlt_solve_dmma: N % 4 == 0");
  constexpr int NW = NT / 32, TM = (N + 7) / 8, TN = (NR + 7) / 8, NB = N / 4;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
#pragma unroll 1
  for (int tj = warp; tj < TN; tj += NW) {
    const int jc = tj * 8 + 2 * q;
    double c[TM][2];
#pragma unroll
    for (int mt = 0; mt < TM; mt++) {
      const int row = g + 8 * mt;
      c[mt][0] = (row < N && jc < NR) ? R.get(row, jc) : 0.0;
      c[mt][1] = (row < N && jc + 1 < NR) ? R.get(row, jc + 1) : 0.0;
    }
    // forward: L Y = R
#pragma unroll
    for (int kb = 0; kb < NB; kb++) {
      const int mt0 = kb >> 1, h = kb & 1, r0 = 4 * kb;
#pragma unroll
      for (int i = 0; i < 3; i++) {                // (a) rows r0 + i + 1 .. r0 + 3 -= L(., r0 + i) y_i
        const int src = ((4 * h + i) << 2) | q;
        const double y0 = __shfl_sync(0xffffffffu, c[mt0][0], src), y1 = __shfl_sync(0xffffffffu, c[mt0][1], src);
        if (g > 4 * h + i && g <= 4 * h + 3) {
          const double l = S[(r0 + g - 4 * h) + (r0 + i) * N];
          c[mt0][0] = fma(-l, y0, c[mt0][0]); c[mt0][1] = fma(-l, y1, c[mt0][1]);
        }
      }
      if (kb + 1 < NB) {                           // (b) rows below -= L(below, block) Y_block
        const int srcb = ((4 * h + q) << 2) | (g >> 1);
        const double b0 = __shfl_sync(0xffffffffu, c[mt0][0], srcb), b1 = __shfl_sync(0xffffffffu, c[mt0][1], srcb);
        const double bv = (g & 1) ? b1 : b0;
#pragma unroll
        for (int mt = mt0; mt < TM; mt++) {
          if (mt == mt0 && h == 1) continue;
          const int row = 8 * mt + g;
          const double a = (row > r0 + 3 && row < N) ? -S[row + (r0 + q) * N] : 0.0;
          dmma_884(c[mt][0], c[mt][1], a, bv);
        }
      }
    }
#pragma unroll
    for (int mt = 0; mt < TM; mt++) {
      const int row = g + 8 * mt;
      if (row < N) { const double d = dinv[row]; c[mt][0] *= d; c[mt][1] *= d; }
    }
    // backward: L^T X = D^{-1} Y
#pragma unroll
    for (int kb = NB - 1; kb >= 0; kb--) {
      const int mt0 = kb >> 1, h = kb & 1, r0 = 4 * kb;
#pragma unroll
      for (int i = 3; i > 0; i--) {                // (a) rows r0 .. r0 + i - 1 -= L(r0 + i, .)^T x_i
        const int src = ((4 * h + i) << 2) | q;
        const double x0 = __shfl_sync(0xffffffffu, c[mt0][0], src), x1 = __shfl_sync(0xffffffffu, c[mt0][1], src);
        if (g >= 4 * h && g < 4 * h + i) {
          const double l = S[(r0 + i) + (r0 + g - 4 * h) * N];
          c[mt0][0] = fma(-l, x0, c[mt0][0]); c[mt0][1] = fma(-l, x1, c[mt0][1]);
        }
      }
      if (kb > 0) {                                // (b) rows above -= L(block, above)^T X_block
        const int srcb = ((4 * h + q) << 2) | (g >> 1);
        const double b0 = __shfl_sync(0xffffffffu, c[mt0][0], srcb), b1 = __shfl_sync(0xffffffffu, c[mt0][1], srcb);
        const double bv = (g & 1) ? b1 : b0;
#pragma unroll
        for (int mt = 0; mt <= mt0; mt++) {
          if (mt == mt0 && h == 0) continue;
          const int row = 8 * mt + g;
          const double a = (row < r0) ? -S[(r0 + q) + row * N] : 0.0;
          dmma_884(c[mt][0], c[mt][1], a, bv);
        }
      }
    }
#pragma unroll
    for (int mt = 0; mt < TM; mt++) {
      const int row = g + 8 * mt;
      if (row < N) {
        if (jc < NR) R.set(row, jc, c[mt][0]);
        if (jc + 1 < NR) R.set(row, jc + 1, c[mt][1]);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// X = E^{-1} entirely in REGISTERS: in-place Gauss-Jordan elimination with row partial
// pivoting (same pivot rule as PartialPivLU: largest |entry| of the column among the rows not
// yet used).  LPM lanes own one matrix (8 for N <= 16: four matrices per warp), lane `sub`
// holds rows sub, sub + LPM, ...  All loops are unrolled so every register index is static;
// per step the pivot row travels through N doubles of shared memory (`srow`) as a broadcast.
// Rows never move: step k records its pivot row in spiv[k] and the row remembers its step,
// and the final store un-permutes:  X(step(r), spiv[c]) = a_r[c].
// Returns false on an exactly zero pivot.
// ------------------------------------------------------------------------------------------
// Reciprocal of a Gauss-Jordan pivot: the 20-bit hardware seed (MUFU.RCP64H) and two Newton steps,
// ~1 ulp, no slow-path branch (a zero / non-finite pivot is caught by the caller's singularity flag).
// The pivot reciprocal sits on the dependent chain of every elimination step: this is shorter than
// __drcp_rn and keeps the unrolled pivot loops free of branches.
__device__ __forceinline__ double rcp_pivot(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

template <int N> struct GjShape {
#ifndef GARB200_GJ_LPM_SMALL
#define GARB200_GJ_LPM_SMALL 8
#endif
  // lanes per matrix: 8 for N <= 16 (four matrices per warp share the per-step pivot search and
  // bookkeeping instructions; rows l and l + 8 per lane), a whole warp otherwise
  static constexpr int LPM = N <= 16 ? GARB200_GJ_LPM_SMALL : 32;
  static constexpr int MPW = 32 / LPM;
  static constexpr int RPL = (N + LPM - 1) / LPM;
  static_assert(N <= 64, "register Gauss-Jordan: N <= 64");
};

template <int N>
__device__ __forceinline__ bool gj_inverse_regs(int sub, unsigned mask, const double *__restrict__ Eg,
                                                double *__restrict__ Xg, bool do_store, double *srow,
                                                int *spiv) {
  constexpr int LPM = GjShape<N>::LPM, RPL = GjShape<N>::RPL;
  double a[RPL][N];
  int sig[RPL];
  bool act[RPL];
  double rscale[RPL];
#pragma unroll
  for (int s = 0; s < RPL; s++) {
    const int row = sub + LPM * s;
    act[s] = row < N;
    sig[s] = 0;
    rscale[s] = 1.0;
#pragma unroll
    for (int j = 0; j < N; j++) a[s][j] = act[s] ? Eg[row + j * N] : 0.0;
  }
  bool ok = true;
  static_assert(N % 2 == 0, "register Gauss-Jordan: even N (128-bit shared-memory accesses)");
  double2 *srow2 = reinterpret_cast<double2 *>(srow);
  // The register file is ROTATED by one column per step (the column being eliminated always
  // sits at index 0, the finished inverse column is appended at index N-1), so the loop body
  // has static register indices without unrolling over k: the whole kernel stays a few hundred
  // instructions and lives in the instruction cache.  After N steps the order is restored.
#pragma unroll 1
  for (int k = 0; k < N; k++) {
    // pivot search: 64-bit key = bits(|a|) with the low 7 bits replaced by (127 - row)
    unsigned long long best = 0ull;
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      const int row = sub + LPM * s;
      unsigned long long key = (unsigned long long)__double_as_longlong(fabs(a[s][0]));
      key = (key & ~0x7Full) | (unsigned long long)(127 - row);
      if (act[s] && key > best) best = key;
    }
    const unsigned hi = (unsigned)(best >> 32), lo = (unsigned)best;
    const unsigned mhi = __reduce_max_sync(mask, hi);
    const unsigned mlo = __reduce_max_sync(mask, (best != 0ull && hi == mhi) ? lo : 0u);
    const int r = 127 - (int)(mlo & 0x7Fu);
    const int rl = r % LPM, rs = r / LPM;
    double pcand = a[0][0];
#pragma unroll
    for (int s = 1; s < RPL; s++) if (rs == s) pcand = a[s][0];
    const double pval = __shfl_sync(mask, pcand, rl, LPM);
    if (pval == 0.0) ok = false;
    const double pinv = rcp_pivot(pval);
    if (sub == rl) {
#pragma unroll
      for (int s = 0; s < RPL; s++)
        if (rs == s) {
#pragma unroll
          for (int j = 0; j < N; j += 2) srow2[j / 2] = make_double2(a[s][j], a[s][j + 1]);
        }
    }
    if (sub == 0) spiv[k] = r;
    __syncwarp(mask);
    double f[RPL];
    bool isp[RPL];
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      isp[s] = (sub + LPM * s) == r;
      f[s] = isp[s] ? 0.0 : a[s][0] * pinv;
    }
#pragma unroll
    for (int j = 0; j < N; j += 2) {
      const double2 bj = srow2[j / 2];
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        if (j > 0) a[s][j - 1] = fma(-f[s], bj.x, a[s][j]);
        a[s][j] = fma(-f[s], bj.y, a[s][j + 1]);
      }
    }
    // A pivot row is NOT scaled by 1/pivot here: it keeps its raw values (new column entry 1)
    // and later steps treat it like any other row, which commutes with the row scaling; the
    // scale factor is applied once in the final store.
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      a[s][N - 1] = isp[s] ? 1.0 : -f[s];
      if (isp[s]) {
        act[s] = false;
        sig[s] = k;
        rscale[s] = pinv;
      }
    }
    __syncwarp(mask);
  }
  __syncwarp(mask);
  if (do_store) {
#pragma unroll
    for (int c = 0; c < N; c++) {
      const int pc = spiv[c];
#pragma unroll
      for (int s = 0; s < RPL; s++)
        if (sub + LPM * s < N) Xg[sig[s] + pc * N] = a[s][c] * rscale[s];
    }
  }
  return ok;
}

// ------------------------------------------------------------------------------------------
// Block Gauss-Jordan inverse on the FP64 tensor pipe: one warp per N x N matrix (N % 4 == 0).
// The (padded) matrix lives in REGISTERS as m8n8k4 accumulator tiles for all N/4 block steps (the
// master copy: the rank-4 updates never reload it) and is MIRRORED row-major in shared memory after
// every update (one double2 store per tile), so that everything indexed by a run-time row -- the
// 4-column panel of the next step, the raw pivot rows of this one, the rows in pivot order at the
// end -- is a plain shared-memory read instead of a compare/select walk over the register tiles
// (that walk was ~30 % of the instructions of the register-only version, which was issue/latency bound).
// Block step kb (columns c0 = 4 kb .. c0 + 3):
//   1. the panel is read row-per-lane from the mirror and eliminated LU style with partial pivoting over
//      ALL not-yet-pivoted rows: pivot ss clears column ss from the columns to its right only and leaves
//      the multiplier in place, for every row alike (a pivot row just stops being a candidate).  With
//      B = A[R, panel] = L_B U (R = the 4 pivot rows) the multipliers l_i of row i give
//      A[i, panel] B^{-1} = l_i L_B^{-1}: W_i = -A[i, panel] B^{-1} is one 4 x 4 unit-triangular back
//      substitution per row.
//   2. W goes to shared memory (A operand), with W = 0 for the pivot rows; the B operand rows are the
//      RAW pivot rows read straight from the mirror, panel columns replaced by I_4 in the fragment.
//   3. one rank-4 update  C += W * pivot_rows  = 1 DMMA per tile (panel columns zeroed first, so they
//      become W), tiles stored to the mirror, the pivot rows' own panel entries patched to e_s there
//      and re-read into the tiles of that tile column.
// The pivot rows themselves are NOT transformed at their block step: they stay raw (panel
// columns e_s) and keep being updated like any other row, which commutes with the pending
// left-multiplication by B^{-1}; all N/4 block transforms are applied at the end (the raw 4 x 4 blocks B
// are kept and inverted then, one lane per block, in the row order the pivoting chose) as one
// block-diagonal DMMA product on the mirror rows taken in pivot order, which is also where the row
// permutation is undone.  The column permutation is undone by the final store:
//   X(step(r), piv[c]) = (B_kb^{-1} C[R_kb, :])(r, c).
// Mirror layout: element (row, col) at Y[row * PLD + (col ^ swz(row))], PLD == 8 (mod 16) so that the
// double2 tile stores of a quarter warp (rows g, g + 1) cover all 32 banks, swz(row) = 4 * bit1(row) so
// that the row-per-lane panel reads of 8 consecutive rows spread over 4 bank groups instead of 2.
// Shared memory per warp: GjDmma<N>::SMEM doubles + GjDmma<N>::SINTS ints.
// ------------------------------------------------------------------------------------------
template <int N> struct GjDmma {
  static_assert(N % 4 == 0 && N <= 64, "block Gauss-Jordan: N % 4 == 0, N <= 64");
  static constexpr int NTL = (N + 7) / 8, NP = NTL * 8, NB = N / 4;
  static constexpr int PLD = (NP % 16 == 8) ? NP : NP + 8;
  static constexpr int WLD = (NP % 16 == 8) ? NP + 4 : NP + 12;   // W panel, column-major: == 12 (mod 16), conflict-free A fragments
  static constexpr int RPL = (NP + 31) / 32;
  static constexpr int PAN = 0, BINV = 4 * WLD, YS = BINV + 16 * (2 * NTL);
  static constexpr int SMEM = YS + N * PLD;     // Y: row-major mirror of the matrix
  static constexpr int SINTS = NP;              // piv[c]
};

template <int N>
__device__ __forceinline__ bool gj_inverse_dmma(int lane, const double *__restrict__ Eg,
                                                double *__restrict__ Xg, bool do_store,
                                                double *sw, int *si) {
  using G = GjDmma<N>;
  constexpr int NTL = G::NTL, NP = G::NP, NB = G::NB, PLD = G::PLD, WLD = G::WLD, RPL = G::RPL;
  double *pan = sw + G::PAN, *sbinv = sw + G::BINV, *Y = sw + G::YS;
  int *spiv = si;
  const int g = lane >> 2, q = lane & 3;
  const int swg = ((g >> 1) & 1) << 2;          // swizzle of this lane's tile rows g + 8 mt
  const int swl = ((lane >> 1) & 1) << 2;       // ... and of its panel rows lane + 32 s
  double *ytile = Y + g * PLD + ((2 * q) ^ swg);   // tile (mt, nt): ytile[8 * mt * PLD + 8 * nt]
  double c[NTL][NTL][2];
#pragma unroll
  for (int mt = 0; mt < NTL; mt++)
#pragma unroll
    for (int nt = 0; nt < NTL; nt++) {
      const int row = g + 8 * mt, col = 8 * nt + 2 * q;
      c[mt][nt][0] = (row < N && col < N) ? Eg[row + col * N] : 0.0;
      c[mt][nt][1] = (row < N && col + 1 < N) ? Eg[row + (col + 1) * N] : 0.0;
      if (row < N) *reinterpret_cast<double2 *>(ytile + 8 * mt * PLD + 8 * nt) = make_double2(c[mt][nt][0], c[mt][nt][1]);
    }
  // block slots past the last block (padding of the final product) are zero
  for (int e = lane; e < 16 * (2 * NTL - NB); e += 32) sbinv[16 * NB + e] = 0.0;
  bool act[RPL];
#pragma unroll
  for (int s = 0; s < RPL; s++) act[s] = lane + 32 * s < N;
  bool ok = true;
  __syncwarp();
#pragma unroll
  for (int kb = 0; kb < NB; kb++) {
    const int ntp = kb >> 1, h = kb & 1, c0 = 4 * kb;
    // 1. the panel, row per lane, from the mirror
    double a[RPL][4];
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      const int row = lane + 32 * s;
      if (row < N) {
        const double *yp = Y + row * PLD + (c0 ^ swl);
        const double2 lo = *reinterpret_cast<const double2 *>(yp), hi = *reinterpret_cast<const double2 *>(yp + 2);
        a[s][0] = lo.x; a[s][1] = lo.y; a[s][2] = hi.x; a[s][3] = hi.y;
      } else {
        a[s][0] = a[s][1] = a[s][2] = a[s][3] = 0.0;
      }
    }
    int rsel[4];
    double lb10 = 0.0, lb20 = 0.0, lb21 = 0.0, lb30 = 0.0, lb31 = 0.0, lb32 = 0.0;
#pragma unroll
    for (int ss = 0; ss < 4; ss++) {
      // pivot search on a 32-bit key: top 25 bits of |a| (exponent + 13 mantissa bits), row in
      // the low 7 bits; candidates closer than 2^-13 relative are interchangeable as pivots
      unsigned best = 0u;
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        const unsigned hi32 = (unsigned)__double2hiint(a[s][ss]) & 0x7fffffffu;
        const unsigned key = (hi32 & ~0x7Fu) | (unsigned)(127 - (lane + 32 * s));
        if (act[s] && key > best) best = key;
      }
      const unsigned mx = __reduce_max_sync(0xffffffffu, best);
      const int r = 127 - (int)(mx & 0x7Fu);
      const int rl = r & 31, rs = r >> 5;
      rsel[ss] = r;
      double pr[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        double v = a[0][j];
#pragma unroll
        for (int s = 1; s < RPL; s++) if (rs == s) v = a[s][j];
        pr[j] = __shfl_sync(0xffffffffu, v, rl);
      }
      if (mx == 0u || pr[ss] == 0.0) ok = false;
      const double pinv = rcp_pivot(pr[ss]);
      // (the broadcast of pivot row ss carries its multipliers for the earlier pivots: row ss of L_B)
      if (ss == 1) lb10 = pr[0];
      if (ss == 2) { lb20 = pr[0]; lb21 = pr[1]; }
      if (ss == 3) { lb30 = pr[0]; lb31 = pr[1]; lb32 = pr[2]; }
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        const double m = a[s][ss] * pinv;
        a[s][ss] = m;
#pragma unroll
        for (int j = ss + 1; j < 4; j++) a[s][j] = fma(-m, pr[j], a[s][j]);
        if ((lane + 32 * s) == r) act[s] = false;
      }
    }
    // W_i = -l_i L_B^{-1}: back substitution with the unit lower 4 x 4 factor; 2. W (A operand) to shared memory
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      const double w3 = -a[s][3];
      const double w2 = -fma(w3, lb32, a[s][2]);
      const double w1 = -fma(w3, lb31, fma(w2, lb21, a[s][1]));
      const double w0 = -fma(w3, lb30, fma(w2, lb20, fma(w1, lb10, a[s][0])));
      const int row = lane + 32 * s;
      if (row < NP) { pan[row] = w0; pan[WLD + row] = w1; pan[2 * WLD + row] = w2; pan[3 * WLD + row] = w3; }
    }
    if (lane < 4) spiv[c0 + lane] = lane == 0 ? rsel[0] : lane == 1 ? rsel[1] : lane == 2 ? rsel[2] : rsel[3];
    // (the pivot rows' own panel entries were patched to e_s in the mirror at the end of the previous
    //  step: bring that tile column of the register copy up to date before it is updated again)
    if (kb > 0) {
      const int ntq = (kb - 1) >> 1, hq = (kb - 1) & 1;
      if ((q >> 1) == hq) {
#pragma unroll
        for (int mt = 0; mt < NTL; mt++) {
          if (g + 8 * mt < N) {
            const double2 v = *reinterpret_cast<const double2 *>(ytile + 8 * mt * PLD + 8 * ntq);
            c[mt][ntq][0] = v.x; c[mt][ntq][1] = v.y;
          }
        }
      }
    }
    __syncwarp();
    if (lane < 16) {
      const int sr = lane >> 2, sc = lane & 3;
      const int r = sr == 0 ? rsel[0] : sr == 1 ? rsel[1] : sr == 2 ? rsel[2] : rsel[3];
      sbinv[16 * kb + lane] = Y[r * PLD + ((c0 + sc) ^ (((r >> 1) & 1) << 2))];   // B(sr, sc) = A[R, panel], raw
      pan[sc * WLD + r] = 0.0;                                                     // W = 0 for the pivot rows
    }
    __syncwarp();
    // 3. rank-4 update, one DMMA per tile.  B rows: the raw pivot rows from the mirror (k = q)
    double af[NTL], bf[NTL];
    const int rq = q == 0 ? rsel[0] : q == 1 ? rsel[1] : q == 2 ? rsel[2] : rsel[3];
    const double *yq = Y + rq * PLD + (g ^ (((rq >> 1) & 1) << 2));
#pragma unroll
    for (int mt = 0; mt < NTL; mt++) af[mt] = pan[q * WLD + g + 8 * mt];
#pragma unroll
    for (int nt = 0; nt < NTL; nt++) bf[nt] = yq[8 * nt];
    if ((g >> 2) == h) bf[ntp] = (g & 3) == q ? 1.0 : 0.0;        // panel columns of the pivot rows: I_4
    if ((q >> 1) == h) {
#pragma unroll
      for (int mt = 0; mt < NTL; mt++) { c[mt][ntp][0] = 0.0; c[mt][ntp][1] = 0.0; }
    }
#pragma unroll
    for (int mt = 0; mt < NTL; mt++)
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) dmma_884(c[mt][nt][0], c[mt][nt][1], af[mt], bf[nt]);
    __syncwarp();     // every lane has read its B rows before the mirror is overwritten
    // (columns >= N of the last tile column are padding: zero in the mirror since the initial store, never
    //  rewritten -- the shared-memory pipe is what bounds this kernel)
#pragma unroll
    for (int mt = 0; mt < NTL; mt++)
#pragma unroll
      for (int nt = 0; nt < NTL; nt++)
        if (g + 8 * mt < N && 8 * nt + 2 * q < N) *reinterpret_cast<double2 *>(ytile + 8 * mt * PLD + 8 * nt) = make_double2(c[mt][nt][0], c[mt][nt][1]);
    __syncwarp();
    if (lane < 16) {
      const int sr = lane >> 2, sc = lane & 3;
      const int r = sr == 0 ? rsel[0] : sr == 1 ? rsel[1] : sr == 2 ? rsel[2] : rsel[3];
      Y[r * PLD + ((c0 + sc) ^ (((r >> 1) & 1) << 2))] = (sr == sc) ? 1.0 : 0.0;
    }
    __syncwarp();
  }
  // B_kb^{-1} for all block steps at once, one lane per 4 x 4 block: Gauss-Jordan without pivoting in the
  // row order the partial pivoting above chose (= the LU order of B, so it is as stable as that was)
  if (lane < NB) {
    double bm[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j2 = 0; j2 < 4; j2++) bm[i][j2] = sbinv[16 * lane + 4 * i + j2];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const double pinv = rcp_pivot(bm[k][k]);
#pragma unroll
      for (int j2 = 0; j2 < 4; j2++) bm[k][j2] = (j2 == k) ? pinv : bm[k][j2] * pinv;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        if (i == k) continue;
        const double m = bm[i][k];
#pragma unroll
        for (int j2 = 0; j2 < 4; j2++) bm[i][j2] = (j2 == k) ? -m * pinv : fma(-m, bm[k][j2], bm[i][j2]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j2 = 0; j2 < 4; j2++) sbinv[16 * lane + 4 * i + j2] = bm[i][j2];
  }
  __syncwarp();
  // X rows = blockdiag(B_kb^{-1}) (mirror rows in pivot order)
  int pc[NTL][2];
#pragma unroll
  for (int nt = 0; nt < NTL; nt++) {
    const int col = 8 * nt + 2 * q;
    pc[nt][0] = col < N ? spiv[col] : 0;
    pc[nt][1] = col + 1 < N ? spiv[col + 1] : 0;
  }
#pragma unroll
  for (int mt = 0; mt < NTL; mt++) {
    double x[NTL][2];
#pragma unroll
    for (int nt = 0; nt < NTL; nt++) x[nt][0] = x[nt][1] = 0.0;
#pragma unroll
    for (int j2 = 0; j2 < 2; j2++) {
      if (2 * mt + j2 < NB) {
        // A(row g, k q) = B_{2mt+j}^{-1}(g - 4j, q) on rows 4j..4j+3, zero elsewhere
        const double av = ((g >> 2) == j2) ? sbinv[16 * (2 * mt + j2) + 4 * (g & 3) + q] : 0.0;
        const int rr = spiv[8 * mt + 4 * j2 + q];                      // the row pivoted at step 8 mt + 4 j + q
        const double *yr = Y + rr * PLD + (g ^ (((rr >> 1) & 1) << 2));
#pragma unroll
        for (int nt = 0; nt < NTL; nt++) dmma_884(x[nt][0], x[nt][1], av, yr[8 * nt]);
      }
    }
    const int row = g + 8 * mt;
    if (do_store && row < N) {
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) {
        const int col = 8 * nt + 2 * q;
        if (col < N) Xg[row + pc[nt][0] * N] = x[nt][0];
        if (col + 1 < N) Xg[row + pc[nt][1] * N] = x[nt][1];
      }
    }
  }
  return ok;
}

// ------------------------------------------------------------------------------------------
// The same block Gauss-Jordan inverse for matrices too large for one warp's registers (nx = 76):
// one CTA of NWARP warps per matrix, the padded matrix ROW-major in shared memory (double2 tile
// accesses), warp 0 pivots the 4-column panel (row-per-lane, partial pivoting over all unpivoted
// rows) while the other CTAs of the SM run their rank-4 DMMA updates; two barriers per block step.
// Deferred block transforms, final block-diagonal product and the un-permuting store are those of
// gj_inverse_dmma; the "rows sorted by pivot step" copy is replaced by an indirect row lookup.
// ------------------------------------------------------------------------------------------
template <int N> struct GjCta {
  static_assert(N % 4 == 0 && N <= 124, "CTA block Gauss-Jordan: N % 4 == 0, N <= 124 (7-bit row key)");
  static constexpr int NTL = (N + 7) / 8, NP = NTL * 8, NB = N / 4;
  static constexpr int LD = (NP % 16 == 8) ? NP : NP + 8;          // == 8 (mod 16): conflict-free double2 tiles
  static constexpr int PLD = (NP % 16 == 8) ? NP + 4 : NP + 12;    // == 12 (mod 16): conflict-free B fragments
  static constexpr int RPL = (NP + 31) / 32;
  static constexpr int PAN = NP * LD, PROW = PAN + NP * 4, BINV = PROW + 4 * PLD;
  static constexpr int SMEM = BINV + 16 * (2 * NTL);
  static constexpr int SINTS = 2 * NP;
};

template <int N, int NWARP>
__device__ __forceinline__ bool gj_inverse_cta(int tid, const double *__restrict__ Eg, double *__restrict__ Xg,
                                               double *sw, int *si) {
  using G = GjCta<N>;
  constexpr int NTL = G::NTL, NP = G::NP, NB = G::NB, LD = G::LD, PLD = G::PLD, RPL = G::RPL, NT = 32 * NWARP;
  double *M = sw, *pan = sw + G::PAN, *prow = sw + G::PROW, *sbinv = sw + G::BINV;
  int *spiv = si, *sstep = si + NP;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
  for (int e = tid; e < NP * LD; e += NT) M[e] = 0.0;
  for (int e = tid; e < 16 * (2 * NTL - NB); e += NT) sbinv[16 * NB + e] = 0.0;
  __syncthreads();
  {   // transposing fill: a warp moves 4 rows x 8 columns at a time (32-byte runs of a column-major
      // column in global memory, a 2-way instead of a 16-way bank conflict on the row-major store)
    constexpr int RB = N / 4, CB = (N + 7) / 8;
    const int r4 = lane & 3, c8 = lane >> 2;
    for (int blk = warp; blk < RB * CB; blk += NWARP) {
      const int row = (blk % RB) * 4 + r4, col = (blk / RB) * 8 + c8;
      if (col < N) M[row * LD + col] = Eg[row + col * N];
    }
  }
  bool act[RPL];
#pragma unroll
  for (int s = 0; s < RPL; s++) act[s] = lane + 32 * s < N;
  int bad = 0;
  __syncthreads();
#pragma unroll 1
  for (int kb = 0; kb < NB; kb++) {
    const int c0 = 4 * kb;
    if (warp == 0) {
      double a[RPL][4];
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        const int row = lane + 32 * s;
        if (row < NP) {
          const double2 lo = *reinterpret_cast<const double2 *>(M + row * LD + c0), hi = *reinterpret_cast<const double2 *>(M + row * LD + c0 + 2);
          a[s][0] = lo.x; a[s][1] = lo.y; a[s][2] = hi.x; a[s][3] = hi.y;
        } else {
          a[s][0] = a[s][1] = a[s][2] = a[s][3] = 0.0;
        }
      }
      int rsel[4];
#pragma unroll
      for (int ss = 0; ss < 4; ss++) {
        unsigned best = 0u;
#pragma unroll
        for (int s = 0; s < RPL; s++) {
          const unsigned hi32 = (unsigned)__double2hiint(a[s][ss]) & 0x7fffffffu;
          const unsigned key = (hi32 & ~0x7Fu) | (unsigned)(127 - (lane + 32 * s));
          if (act[s] && key > best) best = key;
        }
        const unsigned mx = __reduce_max_sync(0xffffffffu, best);
        const int r = 127 - (int)(mx & 0x7Fu);
        const int rl = r & 31, rs = r >> 5;
        rsel[ss] = r;
        double pr[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          double v = a[0][j];
#pragma unroll
          for (int s = 1; s < RPL; s++) if (rs == s) v = a[s][j];
          pr[j] = __shfl_sync(0xffffffffu, v, rl);
        }
        if (mx == 0u || pr[ss] == 0.0) bad = 1;
        const double pinv = rcp_pivot(pr[ss]);
#pragma unroll
        for (int j = 0; j < 4; j++) pr[j] *= pinv;
#pragma unroll
        for (int s = 0; s < RPL; s++) {
          const bool isp = (lane + 32 * s) == r;
          const double mm = a[s][ss];
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (j != ss) a[s][j] = isp ? pr[j] : fma(-mm, pr[j], a[s][j]);
          a[s][ss] = isp ? pinv : -mm * pinv;
          if (isp) act[s] = false;
        }
      }
      // W -> pan (A operand of the update), the matrix' own panel columns -> 0 (they receive W,
      // resp. e_s for the pivot rows, through the identity block of prow)
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        const int row = lane + 32 * s;
        if (row < NP) {
          *reinterpret_cast<double2 *>(pan + row * 4) = make_double2(a[s][0], a[s][1]);
          *reinterpret_cast<double2 *>(pan + row * 4 + 2) = make_double2(a[s][2], a[s][3]);
          *reinterpret_cast<double2 *>(M + row * LD + c0) = make_double2(0.0, 0.0);
          *reinterpret_cast<double2 *>(M + row * LD + c0 + 2) = make_double2(0.0, 0.0);
        }
      }
      if (lane < 4) {
        const int r = lane == 0 ? rsel[0] : lane == 1 ? rsel[1] : lane == 2 ? rsel[2] : rsel[3];
        spiv[c0 + lane] = r;
        sstep[r] = c0 + lane;
      }
      __syncwarp();
      // raw pivot rows -> prow, panel columns of prow = I_4
#pragma unroll
      for (int sr = 0; sr < 4; sr++) {
        const int r = rsel[sr];
        for (int col = lane; col < NP; col += 32)
          prow[sr * PLD + col] = (col >= c0 && col < c0 + 4) ? ((col - c0 == sr) ? 1.0 : 0.0) : M[r * LD + col];
      }
      if (lane < 16) {
        const int sr = lane >> 2, sc = lane & 3;
        const int r = sr == 0 ? rsel[0] : sr == 1 ? rsel[1] : sr == 2 ? rsel[2] : rsel[3];
        sbinv[16 * kb + lane] = pan[r * 4 + sc];       // B^{-1}(sr, sc)
        pan[r * 4 + sc] = 0.0;                         // W = 0 for the pivot rows
        M[r * LD + c0 + sc] = (sr == sc) ? 1.0 : 0.0;
      }
    }
    __syncthreads();
    // rank-4 update: one DMMA per tile, whole 8-row strips per warp
    for (int mt = warp; mt < NTL; mt += NWARP) {
      const double af = pan[(g + 8 * mt) * 4 + q];
      double2 c[NTL];
      double bf[NTL];
      double *Mr = M + (g + 8 * mt) * LD + 2 * q;
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) { bf[nt] = prow[q * PLD + g + 8 * nt]; c[nt] = *reinterpret_cast<const double2 *>(Mr + 8 * nt); }
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) dmma_884(c[nt].x, c[nt].y, af, bf[nt]);
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) *reinterpret_cast<double2 *>(Mr + 8 * nt) = c[nt];
    }
    __syncthreads();
  }
  // X rows = blockdiag(B_kb^{-1}) * (rows in pivot-step order);  X(step(r), piv[c])
  for (int mt = warp; mt < NTL; mt += NWARP) {
    double x[NTL][2];
#pragma unroll
    for (int nt = 0; nt < NTL; nt++) x[nt][0] = x[nt][1] = 0.0;
#pragma unroll
    for (int j = 0; j < 2; j++) {
      if (2 * mt + j < NB) {
        const double av = ((g >> 2) == j) ? sbinv[16 * (2 * mt + j) + 4 * (g & 3) + q] : 0.0;
        const double *Yr = M + spiv[8 * mt + 4 * j + q] * LD + g;
#pragma unroll
        for (int nt = 0; nt < NTL; nt++) dmma_884(x[nt][0], x[nt][1], av, Yr[8 * nt]);
      }
    }
    const int row = g + 8 * mt;
    if (row < N) {
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) {
        const int col = 8 * nt + 2 * q;
        if (col < N) Xg[row + spiv[col] * N] = x[nt][0];
        if (col + 1 < N) Xg[row + spiv[col + 1] * N] = x[nt][1];
      }
    }
  }
  return __syncthreads_or(bad) == 0;
}

// ------------------------------------------------------------------------------------------
// The block Gauss-Jordan of gj_inverse_dmma for matrices too large for ONE warp's registers (nx = 76), with
// the tiles still in REGISTERS: a CTA of NWARP warps per matrix, warp w owns the tile columns
// [w CPW, (w + 1) CPW) as accumulator fragments for all N / 4 block steps; the row-major swizzled mirror in
// shared memory (same layout as gj_inverse_dmma) carries everything indexed by a run-time row.  Per step:
// warp 0 eliminates the 4-column panel (row-per-lane from the mirror, LU style, partial pivoting over all
// unpivoted rows) and publishes W and the pivot rows' indices; after one barrier every warp updates ITS tile
// columns with one rank-4 DMMA per tile (B rows: the raw pivot rows of its own columns from the mirror),
// stores them to the mirror, the owner of the panel's tile column patches the pivot rows' panel entries to e_s
// and re-reads them; one more barrier.  Against gj_inverse_cta (matrix in shared memory: every update loads
// AND stores all of it) this halves the shared-memory traffic, which is what bounds that kernel.
// ------------------------------------------------------------------------------------------
template <int N, int NWARP> struct GjCtaRegs {
  static_assert(N % 4 == 0 && N <= 124, "CTA block Gauss-Jordan: N % 4 == 0, N <= 124 (7-bit row key)");
  static constexpr int NTL = (N + 7) / 8, NP = NTL * 8, NB = N / 4;
  static_assert(NTL % NWARP == 0, "tile columns must divide evenly over the warps");
  static constexpr int CPW = NTL / NWARP;
  static constexpr int PLD = (NP % 16 == 8) ? NP : NP + 8;          // == 8 (mod 16)
  static constexpr int WLD = (NP % 16 == 8) ? NP + 4 : NP + 12;     // == 12 (mod 16)
  static constexpr int RPL = (NP + 31) / 32;
  static constexpr int PAN = 0, BINV = 4 * WLD, YS = BINV + 16 * (2 * NTL);
  static constexpr int SMEM = YS + N * PLD;
  static constexpr int SINTS = NP + 8;                               // piv[c] | rsel[4] | bad
};

template <int N, int NWARP>
__device__ __forceinline__ bool gj_inverse_cta_regs(int tid, const double *__restrict__ Eg, double *__restrict__ Xg,
                                                    double *sw, int *si) {
  using G = GjCtaRegs<N, NWARP>;
  constexpr int NTL = G::NTL, NP = G::NP, NB = G::NB, PLD = G::PLD, WLD = G::WLD, RPL = G::RPL, CPW = G::CPW;
  double *pan = sw + G::PAN, *sbinv = sw + G::BINV, *Y = sw + G::YS;
  int *spiv = si, *srsel = si + NP, *sbad = si + NP + 4;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
  const int swg = ((g >> 1) & 1) << 2, swl = ((lane >> 1) & 1) << 2;
  const int nt0 = warp * CPW;                                        // first tile column of this warp
  double *ytile = Y + g * PLD + ((2 * q) ^ swg);                     // tile (mt, nt): ytile[8 mt PLD + 8 nt]
  double c[NTL][CPW][2];
#pragma unroll
  for (int mt = 0; mt < NTL; mt++)
#pragma unroll
    for (int u = 0; u < CPW; u++) {
      const int row = g + 8 * mt, col = 8 * (nt0 + u) + 2 * q;
      c[mt][u][0] = (row < N && col < N) ? Eg[row + col * N] : 0.0;
      c[mt][u][1] = (row < N && col + 1 < N) ? Eg[row + (col + 1) * N] : 0.0;
      if (row < N) *reinterpret_cast<double2 *>(ytile + 8 * mt * PLD + 8 * (nt0 + u)) = make_double2(c[mt][u][0], c[mt][u][1]);
    }
  for (int e = tid; e < 16 * (2 * NTL - NB); e += 32 * NWARP) sbinv[16 * NB + e] = 0.0;
  if (tid == 0) *sbad = 0;
  bool act[RPL];
#pragma unroll
  for (int s = 0; s < RPL; s++) act[s] = lane + 32 * s < N;
#pragma unroll 1
  for (int kb = 0; kb < NB; kb++) {
    const int ntp = kb >> 1, h = kb & 1, c0 = 4 * kb;
    __syncthreads();                              // mirror current (tiles + the previous step's patch)
    if (warp == 0) {
      double a[RPL][4];
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        const int row = lane + 32 * s;
        if (row < N) {
          const double *yp = Y + row * PLD + (c0 ^ swl);
          const double2 lo = *reinterpret_cast<const double2 *>(yp), hi = *reinterpret_cast<const double2 *>(yp + 2);
          a[s][0] = lo.x; a[s][1] = lo.y; a[s][2] = hi.x; a[s][3] = hi.y;
        } else {
          a[s][0] = a[s][1] = a[s][2] = a[s][3] = 0.0;
        }
      }
      int rsel[4];
      double lb10 = 0.0, lb20 = 0.0, lb21 = 0.0, lb30 = 0.0, lb31 = 0.0, lb32 = 0.0;
#pragma unroll
      for (int ss = 0; ss < 4; ss++) {
        unsigned best = 0u;
#pragma unroll
        for (int s = 0; s < RPL; s++) {
          const unsigned hi32 = (unsigned)__double2hiint(a[s][ss]) & 0x7fffffffu;
          const unsigned key = (hi32 & ~0x7Fu) | (unsigned)(127 - (lane + 32 * s));
          if (act[s] && key > best) best = key;
        }
        const unsigned mx = __reduce_max_sync(0xffffffffu, best);
        const int r = 127 - (int)(mx & 0x7Fu);
        const int rl = r & 31, rs = r >> 5;
        rsel[ss] = r;
        double pr[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          double v = a[0][j];
#pragma unroll
          for (int s = 1; s < RPL; s++) if (rs == s) v = a[s][j];
          pr[j] = __shfl_sync(0xffffffffu, v, rl);
        }
        if ((mx == 0u || pr[ss] == 0.0) && lane == 0) *sbad = 1;
        const double pinv = rcp_pivot(pr[ss]);
        if (ss == 1) lb10 = pr[0];
        if (ss == 2) { lb20 = pr[0]; lb21 = pr[1]; }
        if (ss == 3) { lb30 = pr[0]; lb31 = pr[1]; lb32 = pr[2]; }
#pragma unroll
        for (int s = 0; s < RPL; s++) {
          const double m = a[s][ss] * pinv;
          a[s][ss] = m;
#pragma unroll
          for (int j = ss + 1; j < 4; j++) a[s][j] = fma(-m, pr[j], a[s][j]);
          if ((lane + 32 * s) == r) act[s] = false;
        }
      }
#pragma unroll
      for (int s = 0; s < RPL; s++) {
        const double w3 = -a[s][3];
        const double w2 = -fma(w3, lb32, a[s][2]);
        const double w1 = -fma(w3, lb31, fma(w2, lb21, a[s][1]));
        const double w0 = -fma(w3, lb30, fma(w2, lb20, fma(w1, lb10, a[s][0])));
        const int row = lane + 32 * s;
        if (row < NP) { pan[row] = w0; pan[WLD + row] = w1; pan[2 * WLD + row] = w2; pan[3 * WLD + row] = w3; }
      }
      if (lane < 4) {
        const int r = lane == 0 ? rsel[0] : lane == 1 ? rsel[1] : lane == 2 ? rsel[2] : rsel[3];
        spiv[c0 + lane] = r; srsel[lane] = r;
      }
      __syncwarp();
      if (lane < 16) {
        const int sr = lane >> 2, sc = lane & 3;
        const int r = sr == 0 ? rsel[0] : sr == 1 ? rsel[1] : sr == 2 ? rsel[2] : rsel[3];
        sbinv[16 * kb + lane] = Y[r * PLD + ((c0 + sc) ^ (((r >> 1) & 1) << 2))];   // B(sr, sc), raw
        pan[sc * WLD + r] = 0.0;                                                     // W = 0 for the pivot rows
      }
    }
    __syncthreads();
    // rank-4 update of this warp's tile columns
    const int rq = srsel[q];
    const double *yq = Y + rq * PLD + (g ^ (((rq >> 1) & 1) << 2));
    double bf[CPW];
#pragma unroll
    for (int u = 0; u < CPW; u++) {
      bf[u] = yq[8 * (nt0 + u)];
      if (nt0 + u == ntp && (g >> 2) == h) bf[u] = (g & 3) == q ? 1.0 : 0.0;       // panel columns of the pivot rows: I_4
    }
#pragma unroll
    for (int u = 0; u < CPW; u++)
      if (nt0 + u == ntp && (q >> 1) == h) {
#pragma unroll
        for (int mt = 0; mt < NTL; mt++) { c[mt][u][0] = 0.0; c[mt][u][1] = 0.0; }
      }
#pragma unroll
    for (int mt = 0; mt < NTL; mt++) {
      const double af = pan[q * WLD + g + 8 * mt];
#pragma unroll
      for (int u = 0; u < CPW; u++) dmma_884(c[mt][u][0], c[mt][u][1], af, bf[u]);
    }
    __syncwarp();     // every lane has read its B rows before this warp overwrites its mirror columns
#pragma unroll
    for (int mt = 0; mt < NTL; mt++)
#pragma unroll
      for (int u = 0; u < CPW; u++)
        if (g + 8 * mt < N && 8 * (nt0 + u) + 2 * q < N)
          *reinterpret_cast<double2 *>(ytile + 8 * mt * PLD + 8 * (nt0 + u)) = make_double2(c[mt][u][0], c[mt][u][1]);
    if (ntp >= nt0 && ntp < nt0 + CPW) {          // the owner of the panel's tile column (warp-uniform)
      __syncwarp();
      if (lane < 16) {
        const int sr = lane >> 2, sc = lane & 3;
        const int r = srsel[sr];
        Y[r * PLD + ((c0 + sc) ^ (((r >> 1) & 1) << 2))] = (sr == sc) ? 1.0 : 0.0;
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < CPW; u++)
        if (nt0 + u == ntp && (q >> 1) == h) {
#pragma unroll
          for (int mt = 0; mt < NTL; mt++)
            if (g + 8 * mt < N) {
              const double2 v = *reinterpret_cast<const double2 *>(ytile + 8 * mt * PLD + 8 * ntp);
              c[mt][u][0] = v.x; c[mt][u][1] = v.y;
            }
        }
    }
  }
  __syncthreads();
  // B_kb^{-1} for all block steps, one thread per 4 x 4 block: Gauss-Jordan in the pivoting's row order
  if (tid < NB) {
    double bm[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j2 = 0; j2 < 4; j2++) bm[i][j2] = sbinv[16 * tid + 4 * i + j2];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const double pinv = rcp_pivot(bm[k][k]);
#pragma unroll
      for (int j2 = 0; j2 < 4; j2++) bm[k][j2] = (j2 == k) ? pinv : bm[k][j2] * pinv;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        if (i == k) continue;
        const double m = bm[i][k];
#pragma unroll
        for (int j2 = 0; j2 < 4; j2++) bm[i][j2] = (j2 == k) ? -m * pinv : fma(-m, bm[k][j2], bm[i][j2]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j2 = 0; j2 < 4; j2++) sbinv[16 * tid + 4 * i + j2] = bm[i][j2];
  }
  __syncthreads();
  // X rows = blockdiag(B_kb^{-1}) (mirror rows in pivot order); row tiles over the warps
  for (int mt = warp; mt < NTL; mt += NWARP) {
    double x[NTL][2];
#pragma unroll
    for (int nt = 0; nt < NTL; nt++) x[nt][0] = x[nt][1] = 0.0;
#pragma unroll
    for (int j2 = 0; j2 < 2; j2++) {
      if (2 * mt + j2 < NB) {
        const double av = ((g >> 2) == j2) ? sbinv[16 * (2 * mt + j2) + 4 * (g & 3) + q] : 0.0;
        const int rr = spiv[8 * mt + 4 * j2 + q];
        const double *yr = Y + rr * PLD + (g ^ (((rr >> 1) & 1) << 2));
#pragma unroll
        for (int nt = 0; nt < NTL; nt++) dmma_884(x[nt][0], x[nt][1], av, yr[8 * nt]);
      }
    }
    const int row = g + 8 * mt;
    if (row < N) {
#pragma unroll
      for (int nt = 0; nt < NTL; nt++) {
        const int col = 8 * nt + 2 * q;
        if (col < N) Xg[row + spiv[col] * N] = x[nt][0];
        if (col + 1 < N) Xg[row + spiv[col + 1] * N] = x[nt][1];
      }
    }
  }
  return *sbad == 0;
}

} // namespace garb200
