// device_linalg.cuh — thread-group dense kernels on shared-memory matrices (FP64).
//
// A "group" is the set of NT threads that cooperate on ONE LQ problem: a whole CTA, a
// named-barrier slice of a CTA, or a single warp (small nx: several problems per CTA).
// All matrices are addressed through (pointer, row stride, col stride) so transposes and
// row-major gain blocks are free views.
//
// What each routine restates (numerics follow the reference; summation order does not):
//   lu_inverse     Eigen::PartialPivLU(E).solve(I)      riccati-kernel.hxx:172-174
//   spd_solve      Eigen::LLT(I + mu*Ptilde).solveInPlace riccati-kernel.hxx:185-189,265
//   bk_factor/solve Eigen::BunchKaufman                  core/bunchkaufman.hpp:20-167,389-415,447-516
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace garb200 {

// ------------------------------------------------------------------------------------
// Group
// ------------------------------------------------------------------------------------
template <int NT_> struct Group {
  static constexpr int NT = NT_;
  int tid;     // 0 .. NT-1 inside the group
  int bar_id;  // hardware named barrier (0 == whole CTA) used when NT > 32
  __device__ __forceinline__ void sync() const {
    if constexpr (NT == 32) {
      __syncwarp();
    } else {
      asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(NT) : "memory");
    }
  }
  __device__ __forceinline__ int lane() const { return tid & 31; }
  __device__ __forceinline__ int warp() const { return tid >> 5; }
};

// Barrier over the NT threads that cooperate on ONE problem.  GRP > 1: the CTA holds GRP problems
// (threads [g NT, (g+1) NT) own problem g) and each group has its own hardware named barrier
// 1 + g; barrier 0 (__syncthreads) stays free for aligning the groups with each other.
template <int NT, int GRP = 1> __device__ __forceinline__ void cta_sync() {
  if constexpr (NT == 32) __syncwarp();
  else if constexpr (GRP == 1) __syncthreads();
  else asm volatile("bar.sync %0, %1;" ::"r"(1 + (int)(threadIdx.x / NT)), "r"(NT) : "memory");
}
// group-wide AND of a predicate (the __syncthreads_and of one group)
template <int NT, int GRP = 1> __device__ __forceinline__ int cta_all(int pred) {
  if constexpr (NT == 32) return __all_sync(0xffffffffu, pred);
  else if constexpr (GRP == 1) return __syncthreads_and(pred);
  else {
    int out;
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 q, %1, 0;\n\tbar.red.and.pred p, %2, %3, q;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(out) : "r"(pred), "r"(1 + (int)(threadIdx.x / NT)), "r"(NT) : "memory");
    return out;
  }
}

__device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }

// ------------------------------------------------------------------------------------
// GEMM:  C(i,j) = alpha * sum_k A(i,k) B(k,j) + gin * Cin(i,j)      (Cin may alias C)
// Each thread owns TM x TN register tiles; consecutive threads take consecutive row tiles.
// No internal synchronisation: the caller syncs before (operands ready) and after.
// ------------------------------------------------------------------------------------
template <int TM, int TN, class G>
__device__ __forceinline__ void gemm(const G &g, int M, int N, int K, double alpha,
                                     const double *__restrict__ A, int ars, int acs,
                                     const double *__restrict__ B, int brs, int bcs,
                                     double gin, const double *Cin, int irs, int ics,
                                     double *C, int crs, int ccs) {
  if (M <= 0 || N <= 0) return;
  const int tm = (M + TM - 1) / TM, tn = (N + TN - 1) / TN;
  for (int t = g.tid; t < tm * tn; t += G::NT) {
    const int ti = t % tm, tj = t / tm;
    const int i0 = ti * TM, j0 = tj * TN;
    int ia[TM], jb[TN];
#pragma unroll
    for (int r = 0; r < TM; r++) ia[r] = imin(i0 + r, M - 1) * ars;
#pragma unroll
    for (int c = 0; c < TN; c++) jb[c] = imin(j0 + c, N - 1) * bcs;
    double acc[TM][TN];
#pragma unroll
    for (int r = 0; r < TM; r++)
#pragma unroll
      for (int c = 0; c < TN; c++) acc[r][c] = 0.0;
#pragma unroll 4
    for (int k = 0; k < K; k++) {
      double a[TM], b[TN];
#pragma unroll
      for (int r = 0; r < TM; r++) a[r] = A[ia[r] + k * acs];
#pragma unroll
      for (int c = 0; c < TN; c++) b[c] = B[k * brs + jb[c]];
#pragma unroll
      for (int r = 0; r < TM; r++)
#pragma unroll
        for (int c = 0; c < TN; c++) acc[r][c] = fma(a[r], b[c], acc[r][c]);
    }
#pragma unroll
    for (int r = 0; r < TM; r++)
#pragma unroll
      for (int c = 0; c < TN; c++) {
        const int i = i0 + r, j = j0 + c;
        if (i < M && j < N) {
          double v = alpha * acc[r][c];
          if (Cin) v = fma(gin, Cin[i * irs + j * ics], v);
          C[i * crs + j * ccs] = v;
        }
      }
  }
}

// y(i) = alpha * sum_k A(i,k) x(k) + gin * yin(i)     (one thread per row, tiny sizes)
template <class G>
__device__ __forceinline__ void gemv(const G &g, int M, int K, double alpha,
                                     const double *__restrict__ A, int ars, int acs,
                                     const double *__restrict__ x, int xs, double gin,
                                     const double *yin, int yis, double *y, int ys) {
  for (int i = g.tid; i < M; i += G::NT) {
    double acc0 = 0.0, acc1 = 0.0;
    int k = 0;
    for (; k + 1 < K; k += 2) {
      acc0 = fma(A[i * ars + k * acs], x[k * xs], acc0);
      acc1 = fma(A[i * ars + (k + 1) * acs], x[(k + 1) * xs], acc1);
    }
    if (k < K) acc0 = fma(A[i * ars + k * acs], x[k * xs], acc0);
    double v = alpha * (acc0 + acc1);
    if (yin) v = fma(gin, yin[i * yis], v);
    y[i * ys] = v;
  }
}

// A(j,i) = A(i,j) for i > j  — selfadjointView<Lower> (riccati-kernel.hxx:180,190)
template <class G>
__device__ __forceinline__ void mirror_lower(const G &g, int n, double *A, int ld) {
  for (int e = g.tid; e < n * n; e += G::NT) {
    const int i = e % n, j = e / n;
    if (i > j) A[j + i * ld] = A[i + j * ld];
  }
}

template <class G>
__device__ __forceinline__ void copy_mat(const G &g, int M, int N, const double *S, int srs,
                                         int scs, double scale, double *D, int drs, int dcs) {
  for (int e = g.tid; e < M * N; e += G::NT) {
    const int i = e % M, j = e / M;
    D[i * drs + j * dcs] = scale * S[i * srs + j * scs];
  }
}

// ------------------------------------------------------------------------------------
// Warp arg-max with "first position wins" tie-break (Eigen's maxCoeff(&idx) semantics).
// ------------------------------------------------------------------------------------
__device__ __forceinline__ void warp_argmax(double &v, int &key, int &idx) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double v2 = __shfl_xor_sync(0xffffffffu, v, o);
    const int k2 = __shfl_xor_sync(0xffffffffu, key, o);
    const int i2 = __shfl_xor_sync(0xffffffffu, idx, o);
    if (v2 > v || (v2 == v && k2 < key)) { v = v2; key = k2; idx = i2; }
  }
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ------------------------------------------------------------------------------------
// lu_inverse:  X <- E^{-1} by Gaussian elimination with row partial pivoting applied to the
// augmented system [E | I] (== PartialPivLU(E).solve(I)), then back substitution.
//   E  n x n col-major (ld n), destroyed      Y  n x n col-major scratch
//   X  n x n col-major output                 iw int scratch [3n]
// Rows are never moved: `pos` tracks where an explicit-swap implementation would hold each
// row, so the pivot SEQUENCE (including tie-breaks) is the reference's.
// One group sync per elimination / substitution step.  Returns false on a zero pivot.
// ------------------------------------------------------------------------------------
template <class G>
__device__ __forceinline__ bool lu_inverse(const G &g, int n, double *E, double *Y, double *X,
                                           int *iw) {
  int *pos = iw, *rowat = iw + n, *pivrow = iw + 2 * n;  // pos[row], rowat[position], pivrow[step]
  for (int e = g.tid; e < n * n; e += G::NT) Y[e] = ((e % n) == (e / n)) ? 1.0 : 0.0;
  for (int i = g.tid; i < n; i += G::NT) { pos[i] = i; rowat[i] = i; }
  g.sync();
  const int rb = imin(n, G::NT);
  const int c0 = g.tid / rb, ncg = imax(G::NT / rb, 1);
  const int r0 = (c0 < ncg) ? g.tid % rb : (1 << 28);  // surplus threads own no rows
  bool ok = true;
  for (int k = 0; k < n; k++) {
    // pivot search (every warp redundantly; candidates = rows whose position is >= k)
    double bv = -1.0; int bk = 0x7fffffff, br = 0;
    for (int i = g.lane(); i < n; i += 32) {
      const int p = pos[i];
      if (p >= k) {
        const double v = fabs(E[i + k * n]);
        if (v > bv || (v == bv && p < bk)) { bv = v; bk = p; br = i; }
      }
    }
    warp_argmax(bv, bk, br);
    const int r = br;
    const double piv = E[r + k * n];
    if (bv == 0.0) ok = false;
    const double pinv = 1.0 / piv;
    g.sync();  // everyone has read pos[] before thread 0 edits it
    if (g.tid == 0) {
      const int q = rowat[k], pr = pos[r];
      pos[q] = pr; rowat[pr] = q; pos[r] = k; rowat[k] = r; pivrow[k] = r;
    }
    // eliminate column k from every not-yet-pivoted row (pos > k after the swap <=> pos>=k, i != r)
    for (int i = r0; i < n; i += rb) {
      if (i == r) continue;
      // pos[] may be mid-update by thread 0 only for rows r and q; q's new position is
      // the old position of r (>= k) so its "still active" status is unchanged.
      const int p = pos[i];
      if (p < k) continue;
      const double l = E[i + k * n] * pinv;
      if (l == 0.0) continue;
      for (int j = k + 1 + c0; j < n; j += ncg) E[i + j * n] = fma(-l, E[r + j * n], E[i + j * n]);
      for (int j = c0; j < n; j += ncg) Y[i + j * n] = fma(-l, Y[r + j * n], Y[i + j * n]);
    }
    g.sync();
  }
  // back substitution  U x = y  (column-oriented, one sync per step)
  for (int k = n - 1; k >= 0; k--) {
    const int r = pivrow[k];
    const double uinv = 1.0 / E[r + k * n];
    for (int s = r0; s < k; s += rb) {        // earlier steps' rows
      const int rs = pivrow[s];
      const double u = E[rs + k * n] * uinv;
      if (u == 0.0) continue;
      for (int j = c0; j < n; j += ncg) Y[rs + j * n] = fma(-u, Y[r + j * n], Y[rs + j * n]);
    }
    for (int j = g.tid; j < n; j += G::NT) X[k + j * n] = Y[r + j * n] * uinv;
    g.sync();
  }
  return ok;
}

// ------------------------------------------------------------------------------------
// spd_solve: solve  S Z = RHS  for S symmetric positive definite (lower triangle read,
// destroyed), RHS given as up to three views solved in place.  Square-root-free LDL^T
// elimination fused with the forward substitution, then a column-oriented backward
// substitution: algebraically identical to LLT::solveInPlace.  Returns false if a pivot
// is not positive (== LLT reporting NumericalIssue).
// ------------------------------------------------------------------------------------
struct RhsView { double *p; int rs, cs, ncols; };

template <class G>
__device__ __forceinline__ bool spd_solve(const G &g, int n, double *S, int ld, RhsView v0,
                                          RhsView v1, RhsView v2) {
  const int rb = imin(n, G::NT);
  const int c0 = g.tid / rb, ncg = imax(G::NT / rb, 1);
  const int r0 = (c0 < ncg) ? g.tid % rb : (1 << 28);  // surplus threads own no rows
  bool ok = true;
  RhsView vs[3] = {v0, v1, v2};
  for (int k = 0; k < n; k++) {
    const double d = S[k + k * ld];
    if (!(d > 0.0)) ok = false;
    const double dinv = 1.0 / d;
    for (int i = k + 1 + r0; i < n; i += rb) {
      const double l = S[i + k * ld] * dinv;
      for (int j = k + 1 + c0; j <= i; j += ncg) S[i + j * ld] = fma(-l, S[j + k * ld], S[i + j * ld]);
#pragma unroll
      for (int q = 0; q < 3; q++) {
        double *p = vs[q].p; const int rs = vs[q].rs, cs = vs[q].cs, nc = vs[q].ncols;
        for (int j = c0; j < nc; j += ncg) p[i * rs + j * cs] = fma(-l, p[k * rs + j * cs], p[i * rs + j * cs]);
      }
    }
    g.sync();
  }
  for (int k = n - 1; k >= 0; k--) {
    const double dinv = 1.0 / S[k + k * ld];
    // rows i < k:  b_i -= S(k,i) * x_k   with x_k = b_k * dinv  (recomputed on the fly)
    for (int i = r0; i < k; i += rb) {
      const double l = S[k + i * ld] * dinv;
#pragma unroll
      for (int q = 0; q < 3; q++) {
        double *p = vs[q].p; const int rs = vs[q].rs, cs = vs[q].cs, nc = vs[q].ncols;
        for (int j = c0; j < nc; j += ncg) p[i * rs + j * cs] = fma(-l, p[k * rs + j * cs], p[i * rs + j * cs]);
      }
    }
    g.sync();
    // scale row k (all reads of the unscaled row k are done)
#pragma unroll
    for (int q = 0; q < 3; q++) {
      double *p = vs[q].p; const int rs = vs[q].rs, cs = vs[q].cs, nc = vs[q].ncols;
      for (int j = g.tid; j < nc; j += G::NT) p[k * rs + j * cs] *= dinv;
    }
  }
  g.sync();
  return ok;
}

// ------------------------------------------------------------------------------------
// Bunch-Kaufman LDL^T, executed by ONE warp (lanes own rows); n x n col-major in shared
// memory, lower triangle.  Restates core/bunchkaufman.hpp:20-167 + the post-processing
// of :389-415.  Output: a = unit-lower L + INVERTED diagonal blocks, sub = sub-diagonal
// of the inverted 2x2 blocks, piv = pivot codes (p >= 0: 1x1, row k <-> p; p < 0: 2x2,
// row k+1 <-> -1-p).  Returns false on NumericalIssue (zero pivot column).
// ------------------------------------------------------------------------------------
__device__ __forceinline__ bool bk_factor_warp(int lane, int n, double *a, int ld, int *piv,
                                               double *sub) {
  const double alpha = (1.0 + sqrt(17.0)) / 8.0;
  if (n == 0) return true;
  if (n == 1) {
    const double a00 = a[0];
    __syncwarp();
    if (lane == 0) { piv[0] = 0; sub[0] = 0.0; if (a00 != 0.0) a[0] = 1.0 / a00; }
    __syncwarp();
    return a00 != 0.0;
  }
  int k = 0;
  while (k < n) {
    int kstep = 1;
    const double abs_akk = fabs(a[k + k * ld]);
    // column maximum below the diagonal (first maximum wins)
    double cv = -1.0; int ci = k + 1, ckey;
    for (int i = k + 1 + lane; i < n; i += 32) {
      const double v = fabs(a[i + k * ld]);
      if (v > cv) { cv = v; ci = i; }
    }
    ckey = ci;
    warp_argmax(cv, ckey, ci);
    const double colmax = cv < 0.0 ? 0.0 : cv;
    const int imax = cv < 0.0 ? k + 1 : ci;
    int kp;
    if (fmax(abs_akk, colmax) == 0.0) return false;
    if (abs_akk >= colmax * alpha) {
      kp = k;
    } else {
      double rm = 0.0;
      for (int j = k + lane; j < imax; j += 32) rm = fmax(rm, fabs(a[imax + j * ld]));
      for (int i = imax + 1 + lane; i < n; i += 32) rm = fmax(rm, fabs(a[i + imax * ld]));
      const double rowmax = warp_max(rm);
      if (abs_akk >= (alpha * colmax) * (colmax / rowmax)) kp = k;
      else if (fabs(a[imax + imax * ld]) >= alpha * rowmax) kp = imax;
      else { kp = imax; kstep = 2; }
    }
    const int kk = k + kstep - 1;
    __syncwarp();
    if (kp != kk) {  // symmetric interchange kk <-> kp inside the trailing block
      for (int i = kp + 1 + lane; i < n; i += 32) {
        const double t = a[i + kk * ld]; a[i + kk * ld] = a[i + kp * ld]; a[i + kp * ld] = t;
      }
      for (int j = kk + 1 + lane; j < kp; j += 32) {
        const double t = a[j + kk * ld]; a[j + kk * ld] = a[kp + j * ld]; a[kp + j * ld] = t;
      }
      if (lane == 0) {
        double t = a[kk + kk * ld]; a[kk + kk * ld] = a[kp + kp * ld]; a[kp + kp * ld] = t;
        if (kstep == 2) { t = a[k + 1 + k * ld]; a[k + 1 + k * ld] = a[kp + k * ld]; a[kp + k * ld] = t; }
      }
      __syncwarp();
    }
    if (kstep == 1) {
      const double d11 = 1.0 / a[k + k * ld];
      for (int i = k + 1 + lane; i < n; i += 32) {
        const double xi = a[i + k * ld];
        for (int j = k + 1; j <= i; j++) {
          const double d11xj = a[j + k * ld] * d11;
          a[i + j * ld] -= d11xj * xi;
        }
      }
      __syncwarp();
      for (int i = k + 1 + lane; i < n; i += 32) a[i + k * ld] *= d11;
      if (lane == 0) { a[k + k * ld] = d11; piv[k] = kp; }
    } else {
      const double a21 = a[k + 1 + k * ld];
      const double d21_abs = fabs(a21);
      const double d21_inv = 1.0 / d21_abs;
      const double d11 = d21_inv * a[k + 1 + (k + 1) * ld];
      const double d22 = d21_inv * a[k + k * ld];
      const double t = 1.0 / ((d11 * d22) - 1.0);
      const double d = t * d21_inv;
      const double d21 = a21 * d21_inv;
      for (int i = k + 2 + lane; i < n; i += 32) {
        const double aik = a[i + k * ld], aik1 = a[i + (k + 1) * ld];
        for (int j = k + 2; j <= i; j++) {
          const double ajk = a[j + k * ld], ajk1 = a[j + (k + 1) * ld];
          const double wk = ((ajk * d11) - (ajk1 * d21)) * d;
          const double wkp1 = ((ajk1 * d22) - (ajk * d21)) * d;
          a[i + j * ld] -= aik * wk + aik1 * wkp1;
        }
      }
      __syncwarp();
      for (int j = k + 2 + lane; j < n; j += 32) {
        const double ajk = a[j + k * ld], ajk1 = a[j + (k + 1) * ld];
        a[j + k * ld] = ((ajk * d11) - (ajk1 * d21)) * d;
        a[j + (k + 1) * ld] = ((ajk1 * d22) - (ajk * d21)) * d;
      }
      if (lane == 0) {
        a[k + k * ld] = d11 * d;
        a[k + 1 + k * ld] = -d21 * d;
        a[k + 1 + (k + 1) * ld] = d22 * d;
        piv[k] = -1 - kp; piv[k + 1] = -1 - kp;
      }
    }
    __syncwarp();
    k += kstep;
  }
  // split off the sub-diagonal of the 2x2 blocks (bunchkaufman.hpp:391-402)
  if (lane == 0) {
    for (int q = 0; q < n;) {
      if (piv[q] < 0) { sub[q] = a[q + 1 + q * ld]; sub[q + 1] = 0.0; a[q + 1 + q * ld] = 0.0; q += 2; }
      else { sub[q] = 0.0; q += 1; }
    }
  }
  __syncwarp();
  // apply later interchanges to the already computed columns of L (:404-415)
  for (int q = 0; q < n;) {
    int p = piv[q];
    int row, step;
    if (p < 0) { p = -1 - p; row = q + 1; step = 2; } else { row = q; step = 1; }
    if (p != row)
      for (int j = lane; j < q; j += 32) {
        const double t = a[row + j * ld]; a[row + j * ld] = a[p + j * ld]; a[p + j * ld] = t;
      }
    __syncwarp();
    q += step;
  }
  return true;
}

// ------------------------------------------------------------------------------------------
// Bunch-Kaufman LDL^T with the trailing update spread over a whole group of NT threads
// (the initial-stage KKT matrix is (nx + nc0)^2: 72 x 72 for nx = 36).  Same pivoting rules and
// output format as bk_factor_warp.  Per pivot: warp 0 searches, decides, interchanges and
// snapshots the pivot column(s) into wk; one barrier; every thread updates its share of the
// trailing lower triangle; one barrier.   wk: 2n + 4 doubles, ctl: 4 ints of scratch.
// ------------------------------------------------------------------------------------------
template <int NT, int GRP = 1>
__device__ __forceinline__ bool bk_factor_cta(int tid, int n, double *a, int ld, int *piv, double *sub,
                                              double *wk, int *ctl) {
  const double alpha = 0.6403882032022076;   // (1 + sqrt(17)) / 8
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int TY = NT / 32;
  auto gsync = [&]() { cta_sync<NT, GRP>(); };
  if (n == 0) return true;
  int k = 0;
  bool ok = true;
  while (k < n) {
    if (warp == 0) {
      int kstep = 1, kp = k, fail = 0;
      const double abs_akk = fabs(a[k + k * ld]);
      double cv = -1.0; int ci = k + 1, ckey;
      for (int i = k + 1 + lane; i < n; i += 32) {
        const double v = fabs(a[i + k * ld]);
        if (v > cv) { cv = v; ci = i; }
      }
      ckey = ci;
      warp_argmax(cv, ckey, ci);
      const double colmax = cv < 0.0 ? 0.0 : cv;
      const int imax = cv < 0.0 ? k + 1 : ci;
      if (fmax(abs_akk, colmax) == 0.0) fail = 1;
      else if (k + 1 >= n || abs_akk >= colmax * alpha) kp = k;
      else {
        double rm = 0.0;
        for (int j = k + lane; j < imax; j += 32) rm = fmax(rm, fabs(a[imax + j * ld]));
        for (int i = imax + 1 + lane; i < n; i += 32) rm = fmax(rm, fabs(a[i + imax * ld]));
        const double rowmax = warp_max(rm);
        if (abs_akk >= (alpha * colmax) * (colmax / rowmax)) kp = k;
        else if (fabs(a[imax + imax * ld]) >= alpha * rowmax) kp = imax;
        else { kp = imax; kstep = 2; }
      }
      const int kk = k + kstep - 1;
      __syncwarp();
      if (!fail && kp != kk) {
        for (int i = kp + 1 + lane; i < n; i += 32) {
          const double t = a[i + kk * ld]; a[i + kk * ld] = a[i + kp * ld]; a[i + kp * ld] = t;
        }
        for (int j = kk + 1 + lane; j < kp; j += 32) {
          const double t = a[j + kk * ld]; a[j + kk * ld] = a[kp + j * ld]; a[kp + j * ld] = t;
        }
        if (lane == 0) {
          double t = a[kk + kk * ld]; a[kk + kk * ld] = a[kp + kp * ld]; a[kp + kp * ld] = t;
          if (kstep == 2) { t = a[k + 1 + k * ld]; a[k + 1 + k * ld] = a[kp + k * ld]; a[kp + k * ld] = t; }
        }
        __syncwarp();
      }
      if (!fail) {
        if (kstep == 1) {
          for (int i = k + 1 + lane; i < n; i += 32) wk[i] = a[i + k * ld];
          if (lane == 0) {
            const double d11 = 1.0 / a[k + k * ld];
            wk[2 * n] = d11; a[k + k * ld] = d11; piv[k] = kp;
          }
        } else {
          for (int i = k + 2 + lane; i < n; i += 32) { wk[i] = a[i + k * ld]; wk[n + i] = a[i + (k + 1) * ld]; }
          if (lane == 0) {
            const double a21 = a[k + 1 + k * ld];
            const double d21_abs = fabs(a21);
            const double d21_inv = 1.0 / d21_abs;
            const double d11 = d21_inv * a[k + 1 + (k + 1) * ld];
            const double d22 = d21_inv * a[k + k * ld];
            const double t = 1.0 / ((d11 * d22) - 1.0);
            const double d = t * d21_inv;
            const double d21 = a21 * d21_inv;
            wk[2 * n] = d11; wk[2 * n + 1] = d22; wk[2 * n + 2] = d21; wk[2 * n + 3] = d;
            a[k + k * ld] = d11 * d; a[k + 1 + k * ld] = -d21 * d; a[k + 1 + (k + 1) * ld] = d22 * d;
            piv[k] = -1 - kp; piv[k + 1] = -1 - kp;
          }
        }
      }
      if (lane == 0) { ctl[0] = kstep; ctl[1] = fail; }
    }
    gsync();
    const int kstep = ctl[0];
    if (ctl[1]) { ok = false; break; }
    const int tx = lane, ty = warp;
    if (kstep == 1) {
      const double d11 = wk[2 * n];
      for (int j = k + 1 + ty; j < n; j += TY) {
        const double d11xj = wk[j] * d11;
        for (int i = j + tx; i < n; i += 32) a[i + j * ld] -= d11xj * wk[i];
      }
      for (int i = k + 1 + tid; i < n; i += NT) a[i + k * ld] = wk[i] * d11;
    } else {
      const double d11 = wk[2 * n], d22 = wk[2 * n + 1], d21 = wk[2 * n + 2], d = wk[2 * n + 3];
      for (int j = k + 2 + ty; j < n; j += TY) {
        const double ajk = wk[j], ajk1 = wk[n + j];
        const double w0 = ((ajk * d11) - (ajk1 * d21)) * d;
        const double w1 = ((ajk1 * d22) - (ajk * d21)) * d;
        for (int i = j + tx; i < n; i += 32) a[i + j * ld] -= wk[i] * w0 + wk[n + i] * w1;
      }
      for (int j = k + 2 + tid; j < n; j += NT) {
        const double ajk = wk[j], ajk1 = wk[n + j];
        a[j + k * ld] = ((ajk * d11) - (ajk1 * d21)) * d;
        a[j + (k + 1) * ld] = ((ajk1 * d22) - (ajk * d21)) * d;
      }
    }
    gsync();
    k += kstep;
  }
  if (!ok) return false;
  if (warp == 0) {
    if (lane == 0) {
      for (int q = 0; q < n;) {
        if (piv[q] < 0) { sub[q] = a[q + 1 + q * ld]; sub[q + 1] = 0.0; a[q + 1 + q * ld] = 0.0; q += 2; }
        else { sub[q] = 0.0; q += 1; }
      }
    }
    __syncwarp();
    for (int q = 0; q < n;) {
      int pp = piv[q];
      int row, step;
      if (pp < 0) { pp = -1 - pp; row = q + 1; step = 2; } else { row = q; step = 1; }
      if (pp != row)
        for (int j = lane; j < q; j += 32) {
          const double t = a[row + j * ld]; a[row + j * ld] = a[pp + j * ld]; a[pp + j * ld] = t;
        }
      __syncwarp();
      q += step;
    }
  }
  gsync();
  return true;
}

// ------------------------------------------------------------------------------------------
// Bunch-Kaufman solve of ONE right-hand side by one warp (bunchkaufman.hpp:447-516).
// x lives in shared memory on entry/exit; in between every lane keeps its entries
// (lane, lane+32, ...; n <= 32*MAXS) in registers and the pivot element travels by shuffle,
// so a triangular sweep costs one shuffle + one fused multiply-add per step instead of a
// shared-memory round trip.  `a` should have an odd leading dimension (row AND column
// accesses are then bank-conflict free).  role: n ints of scratch.
// ------------------------------------------------------------------------------------------
template <int MAXS>
__device__ __forceinline__ void bk_solve_vec_warp(int lane, int n, const double *a, int ld,
                                                  const int *piv, const double *sub, double *x,
                                                  int *role) {
  if (lane == 0) {
    for (int k = 0; k < n;) {       // P x, and the role of every index inside the block diagonal
      int p = piv[k];
      if (p < 0) {
        p = -1 - p;
        role[k] = 1; role[k + 1] = 2;
        if (p != k + 1) { const double t = x[k + 1]; x[k + 1] = x[p]; x[p] = t; }
        k += 2;
      } else {
        role[k] = 0;
        if (p != k) { const double t = x[k]; x[k] = x[p]; x[p] = t; }
        k += 1;
      }
    }
  }
  __syncwarp();
  double xr[MAXS];
#pragma unroll
  for (int s = 0; s < MAXS; s++) { const int i = lane + 32 * s; xr[s] = i < n ? x[i] : 0.0; }
  auto bcast = [&](int k) {
    double v = xr[0];
#pragma unroll
    for (int s = 1; s < MAXS; s++) if ((k >> 5) == s) v = xr[s];
    return __shfl_sync(0xffffffffu, v, k & 31);
  };
  for (int k = 0; k < n; k++) {            // unit lower
    const double xk = bcast(k);
#pragma unroll
    for (int s = 0; s < MAXS; s++) { const int i = lane + 32 * s; if (i > k && i < n) xr[s] = fma(-a[i + k * ld], xk, xr[s]); }
  }
#pragma unroll
  for (int s = 0; s < MAXS; s++) { const int i = lane + 32 * s; if (i < n) x[i] = xr[s]; }
  __syncwarp();
#pragma unroll
  for (int s = 0; s < MAXS; s++) {         // block diagonal (stored inverted)
    const int i = lane + 32 * s;
    if (i < n) {
      const int r = role[i];
      if (r == 0) xr[s] = x[i] * a[i + i * ld];
      else if (r == 1) xr[s] = x[i] * a[i + i * ld] + x[i + 1] * sub[i];
      else xr[s] = x[i] * a[i + i * ld] + x[i - 1] * sub[i - 1];
    }
  }
  for (int k = n - 1; k >= 0; k--) {       // unit upper (L^T), column oriented: row k of L
    const double xk = bcast(k);
#pragma unroll
    for (int s = 0; s < MAXS; s++) { const int j = lane + 32 * s; if (j < k) xr[s] = fma(-a[k + j * ld], xk, xr[s]); }
  }
  __syncwarp();
#pragma unroll
  for (int s = 0; s < MAXS; s++) { const int i = lane + 32 * s; if (i < n) x[i] = xr[s]; }
  __syncwarp();
  if (lane == 0) {
    for (int k = n; k > 0;) {              // undo the interchanges
      k -= 1;
      int p = piv[k];
      if (p < 0) { p = -1 - p; if (p != k) { const double t = x[k]; x[k] = x[p]; x[p] = t; } k -= 1; }
      else if (p != k) { const double t = x[k]; x[k] = x[p]; x[p] = t; }
    }
  }
  __syncwarp();
}

// One thread solves one right-hand-side column in place: x(i) = xp[i*xs].
// Restates bunch_kaufman_solve_in_place, core/bunchkaufman.hpp:447-516.
__device__ __forceinline__ void bk_solve_column(int n, const double *__restrict__ a, int ld,
                                                const int *__restrict__ piv,
                                                const double *__restrict__ sub, double *xp,
                                                int xs) {
#define GX(i) xp[(i) * xs]
  for (int k = 0; k < n;) {
    int p = piv[k];
    if (p < 0) { p = -1 - p; const double t = GX(k + 1); GX(k + 1) = GX(p); GX(p) = t; k += 2; }
    else { const double t = GX(k); GX(k) = GX(p); GX(p) = t; k += 1; }
  }
  for (int k = 0; k < n; k++) {          // unit-lower
    const double xk = GX(k);
    for (int i = k + 1; i < n; i++) GX(i) = fma(-a[i + k * ld], xk, GX(i));
  }
  for (int k = 0; k < n;) {              // block diagonal (stored inverted)
    if (piv[k] < 0) {
      const double akp1k = sub[k], ak = a[k + k * ld], akp1 = a[k + 1 + (k + 1) * ld];
      const double xk = GX(k), xkp1 = GX(k + 1);
      GX(k) = xk * ak + xkp1 * akp1k;
      GX(k + 1) = xkp1 * akp1 + xk * akp1k;
      k += 2;
    } else { GX(k) *= a[k + k * ld]; k += 1; }
  }
  for (int k = n - 1; k >= 0; k--) {     // unit-upper (L^T)
    double s = GX(k);
    for (int i = k + 1; i < n; i++) s = fma(-a[i + k * ld], GX(i), s);
    GX(k) = s;
  }
  for (int k = n; k > 0;) {              // undo interchanges
    k -= 1;
    int p = piv[k];
    if (p < 0) { p = -1 - p; const double t = GX(k); GX(k) = GX(p); GX(p) = t; k -= 1; }
    else { const double t = GX(k); GX(k) = GX(p); GX(p) = t; }
  }
#undef GX
}

// ------------------------------------------------------------------------------------
// 1-D TMA bulk copies + mbarrier (sm_90+ PTX; SASS: UBLKCP / SYNCS)
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {}
}
// global -> shared, completion signalled on `bar` (bytes % 16 == 0, 16-byte aligned)
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                            uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// global -> L2 only (no completion tracking)
__device__ __forceinline__ void l2_prefetch(const void *src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
// shared -> global (bulk async-group completion)
__device__ __forceinline__ void tma_store_1d(void *dst_gmem, const void *src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

} // namespace garb200
