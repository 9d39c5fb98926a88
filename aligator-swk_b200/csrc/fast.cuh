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
#include "device_linalg.cuh"

namespace garb200 {

__device__ __forceinline__ void dmma_884(double &c0, double &c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c0), "+d"(c1)
      : "d"(a), "d"(b));
}

/// Strided matrix view with compile-time strides: element (i,j) at p[i*RS + j*CS].
template <int RS, int CS> struct MV {
  double *p;
  __device__ __forceinline__ double operator()(int i, int j) const { return p[i * RS + j * CS]; }
};

// ------------------------------------------------------------------------------------------
// C(i,j) = cin(i,j) + sum_k a(i,k) b(k,j),  M x N x K, distributed over NW warps.
// a / b / cin / store are functors (lambdas): a(i,k), b(k,j), cin(i,j) -> double,
// store(i,j,v).  Indices handed to a/b/cin are always clamped into range; padded rows /
// columns of the 8x8 tiles compute garbage that is never stored.  K is zero-filled to a
// multiple of 4.  TB = number of 8-column tiles that share one A fragment.
// No barrier inside: the caller synchronises before (operands ready) and after.
// ------------------------------------------------------------------------------------------
template <int M, int N, int K, int NW, int TB = 1, class FA, class FB, class FC, class FS>
__device__ __forceinline__ void mma_gemm(int warp, int lane, FA a, FB b, FC cin, FS store) {
  if constexpr (M <= 0 || N <= 0) return;
  constexpr int TM = (M + 7) / 8, TN = (N + 7) / 8, TNB = (TN + TB - 1) / TB;
  constexpr int KS = (K + 3) / 4;
  const int g = lane >> 2, q = lane & 3;
  for (int t = warp; t < TM * TNB; t += NW) {
    const int ti = t % TM, tjb = t / TM;
    const int ia = min(ti * 8 + g, M - 1);      // A row of this lane
    const int ic = ti * 8 + g;                  // C row
    double acc[TB][2];
    int jb[TB];
#pragma unroll
    for (int u = 0; u < TB; u++) {
      const int tj = tjb * TB + u;
      jb[u] = min(tj * 8 + g, N - 1);           // B column of this lane
      const int jc = tj * 8 + 2 * q;            // C columns jc, jc+1
      acc[u][0] = cin(min(ic, M - 1), min(jc, N - 1));
      acc[u][1] = cin(min(ic, M - 1), min(jc + 1, N - 1));
    }
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      const int k = ks * 4 + q;
      double av;
      if constexpr (K % 4 == 0) av = a(ia, k);
      else av = (k < K) ? a(ia, min(k, K - 1)) : 0.0;
#pragma unroll
      for (int u = 0; u < TB; u++) {
        if (TB > 1 && (tjb * TB + u) >= TN) continue;
        double bv;
        if constexpr (K % 4 == 0) bv = b(k, jb[u]);
        else bv = (k < K) ? b(min(k, K - 1), jb[u]) : 0.0;
        dmma_884(acc[u][0], acc[u][1], av, bv);
      }
    }
#pragma unroll
    for (int u = 0; u < TB; u++) {
      const int tj = tjb * TB + u;
      const int jc = tj * 8 + 2 * q;
      if (tj < TN && ic < M) {
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

template <int NT> __device__ __forceinline__ void cta_sync() {
  if constexpr (NT == 32) __syncwarp();
  else __syncthreads();
}

// A(j,i) = A(i,j), i > j (col-major, ld = N): selfadjointView<Lower>
template <int N, int NT> __device__ __forceinline__ void mirror_lower_ct(int tid, double *A) {
  for (int e = tid; e < N * N; e += NT) {
    const int i = e % N, j = e / N;
    if (i > j) A[j + i * N] = A[i + j * N];
  }
}

// ------------------------------------------------------------------------------------------
// Solve S X = R in place for SPD S (lower triangle read and destroyed, col-major ld N) and
// R an N x NR col-major block (ld N).  Square-root-free LDL^T elimination fused with the
// forward substitution, then a column-oriented backward substitution with the reciprocal
// pivots kept in `dinv` (N doubles of scratch).  One barrier per pivot in each phase.
// Thread (r0, c0): row slot r0 = tid % N, column group c0 = tid / N (CG = NT / N groups).
// Returns false when a pivot is not positive (LLT would report NumericalIssue).
// ------------------------------------------------------------------------------------------
template <int N, int NR, int NT>
__device__ __forceinline__ bool ldlt_solve(int tid, double *S, double *R, double *dinv) {
  constexpr int CG = (NT / N) > 0 ? (NT / N) : 1;
  static_assert(NT >= N, "ldlt_solve needs at least N threads");
  const int c0 = tid / N;
  const int r0 = (c0 < CG) ? tid - c0 * N : (1 << 28);
  bool ok = true;
#pragma unroll 1
  for (int k = 0; k < N; k++) {
    const double d = S[k + k * N];
    if (!(d > 0.0)) ok = false;
    const double di = 1.0 / d;
    const int i = k + 1 + r0;
    if (i < N) {
      const double l = S[i + k * N] * di;
      for (int j = k + 1 + c0; j <= i; j += CG) S[i + j * N] = fma(-l, S[j + k * N], S[i + j * N]);
#pragma unroll 4
      for (int j = c0; j < NR; j += CG) R[i + j * N] = fma(-l, R[k + j * N], R[i + j * N]);
    }
    if (tid == 0) dinv[k] = di;
    cta_sync<NT>();
  }
#pragma unroll 1
  for (int k = N - 1; k >= 0; k--) {
    const double di = dinv[k];
    const int i = r0;                       // rows i < k
    if (i < k) {
      const double l = S[k + i * N] * di;
#pragma unroll 4
      for (int j = c0; j < NR; j += CG) R[i + j * N] = fma(-l, R[k + j * N], R[i + j * N]);
    }
    cta_sync<NT>();
    for (int j = tid; j < NR; j += NT) R[k + j * N] *= di;   // row k is final, nobody reads it again unscaled
  }
  cta_sync<NT>();
  return ok;
}

// ------------------------------------------------------------------------------------------
// X = E^{-1} by LU with row partial pivoting, executed by ONE warp on a row-major work
// array W[N][LDW] (LDW = 2N+1, odd => lane-private rows are bank-conflict free).
// Lanes own rows (lane, lane+32, ...); rows never move, `piv[k]` records the pivot row of
// step k.  Phase 1 factors E in place (multipliers stored below the pivots), phases 2/3 run
// the two triangular solves with ONE OUTPUT COLUMN PER LANE, so they need no barrier.
// Xg: col-major N x N output in global memory.  Returns false on an exactly zero pivot.
// Tie-break on equal |pivot candidates|: lowest row index (Eigen: lowest position).
// ------------------------------------------------------------------------------------------
template <int N>
__device__ __forceinline__ bool lu_inverse_warp(int lane, const double *__restrict__ Eg, double *W,
                                                int *piv, double *__restrict__ Xg) {
  constexpr int LDW = 2 * N + 1;
  constexpr int RPL = (N + 31) / 32;     // rows per lane
  // load E (col-major in global: consecutive lanes -> consecutive addresses)
  for (int j = 0; j < N; j++)
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      const int i = lane + 32 * s;
      if (i < N) W[i * LDW + j] = Eg[i + j * N];
    }
  unsigned active = 0;   // bit s: own row lane+32s not yet used as a pivot
#pragma unroll
  for (int s = 0; s < RPL; s++) if (lane + 32 * s < N) active |= 1u << s;
  __syncwarp();
  bool ok = true;
#pragma unroll 1
  for (int k = 0; k < N; k++) {
    // pivot search: max |W[i][k]| over active rows; 64-bit key = |value| bits, low 6 bits <- ~row
    unsigned long long best = 0ull;
    bool any = false;
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      const int i = lane + 32 * s;
      if ((active >> s) & 1u) {
        unsigned long long key = (unsigned long long)__double_as_longlong(fabs(W[i * LDW + k]));
        key = (key & ~0x7Full) | (unsigned long long)(127 - i);   // N <= 128
        if (!any || key > best) best = key;
        any = true;
      }
    }
    const unsigned hi = any ? (unsigned)(best >> 32) : 0u, lo = (unsigned)best;
    const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
    const unsigned mlo = __reduce_max_sync(0xffffffffu, (any && hi == mhi) ? lo : 0u);
    const int r = 127 - (int)(mlo & 0x7Fu);
    const double pv = W[r * LDW + k];
    if (pv == 0.0) ok = false;
    const double pinv = 1.0 / pv;
    if (lane == 0) piv[k] = r;
#pragma unroll
    for (int s = 0; s < RPL; s++) {
      const int i = lane + 32 * s;
      if (i == r) active &= ~(1u << s);
      else if ((active >> s) & 1u) {
        double *row = W + i * LDW;
        const double *prow = W + r * LDW;
        const double l = row[k] * pinv;
        row[k] = l;
        for (int j = k + 1; j < N; j++) row[j] = fma(-l, prow[j], row[j]);
      }
    }
    __syncwarp();
  }
  // phase 2: Y = L^{-1} P  (column j per lane): y_{r_k} = [r_k == j] - sum_{s<k} l(r_k, s) y_{r_s}
  // phase 3: X = U^{-1} Y               : x_k = (y_{r_k} - sum_{s>k} u(r_k, s) x_s) / u(r_k, k)
  for (int j = lane; j < N; j += 32) {
    double *col = W + N + j;               // element of row i: col[i * LDW]
#pragma unroll 1
    for (int k = 0; k < N; k++) {
      const int rk = piv[k];
      const double *row = W + rk * LDW;
      double acc = (rk == j) ? 1.0 : 0.0;
      for (int s = 0; s < k; s++) acc = fma(-row[s], col[piv[s] * LDW], acc);
      col[rk * LDW] = acc;
    }
#pragma unroll 1
    for (int k = N - 1; k >= 0; k--) {
      const int rk = piv[k];
      const double *row = W + rk * LDW;
      double acc = col[rk * LDW];
      for (int s = k + 1; s < N; s++) acc = fma(-row[s], col[piv[s] * LDW], acc);
      col[rk * LDW] = acc / row[k];
    }
  }
  __syncwarp();
  // X(k, j) = W[piv[k]][N + j]  -> col-major global, consecutive lanes = consecutive k
  for (int j = 0; j < N; j++)
    for (int k = lane; k < N; k += 32) Xg[k + j * N] = W[piv[k] * LDW + N + j];
  return ok;
}

} // namespace garb200
