// consumer.cu — the steps immediately BEFORE and AFTER the LQ solve, on the device,
// so a batched ProxDDP-style caller never stages problems or trajectories through the host:
//   pack_from_device_kernel  SolverProxDDPTpl::updateLQSubproblem      solver-proxddp.hxx:802-873
//                            (knot.A = Jx, knot.B = Ju, knot.E = Jy, knot.Q = Lxx + preg I, ...) from batched,
//                            strided DEVICE arrays of the producer into the packed resident records
//   linear_step_kernel       tryLinearStep on vector spaces             solver-proxddp.hxx:104-148
//                            trial = iterate + alpha * direction for xs, us, vs, lbdas in one pass
//   rollout_kernel           the linear part of tryNonlinearRollout     solver-proxddp.hxx:372-391, 431-446
//                            dus = alpha kff + K dxs, dvs = alpha zff + Z dxs, dlams+ = alpha lff + L dxs,
//                            gains records read in place
#include "device_linalg.cuh"
#include "kernels.cuh"

namespace garb200 {

// element (b, t, i, j) of a producer array
__device__ __forceinline__ double fld(const garb200_device_field &F, int b, int t, int i, int j) {
  return F.ptr ? __ldg(F.ptr + (long long)b * F.stride_b + (long long)t * F.stride_t + i + (long long)j * F.ld) : 0.0;
}

// One CTA per (problem, knot): every block of the packed record from the producer's arrays.
// Blocks are stored compactly with the knot's OWN (nu_t, nc_t) as in garb200_layout.h.
__global__ void __launch_bounds__(128) pack_from_device_kernel(const DevParams p, const garb200_device_problem_view v,
                                                               double preg) {
  constexpr int NT = 128;
  const int T = p.T, n = p.nx, h = p.nth;
  const int b = blockIdx.x / (T + 1), t = blockIdx.x % (T + 1), tid = threadIdx.x;
  const int m = p.nu_t[t], c = p.nc_t[t];
  const garb200_knot_layout &KL = p.KL;
  double *rec = p.knots + ((size_t)b * (T + 1) + phys_t(p, t)) * KL.stride;
  for (int e = tid; e < KL.stride; e += NT) rec[e] = 0.0;
  __syncthreads();
  auto blk = [&](int off, int rows, int cols, const garb200_device_field &F, double diag) {
    for (int e = tid; e < rows * cols; e += NT) {
      const int i = e % rows, j = e / rows;
      rec[off + e] = fld(F, b, t, i, j) + (i == j ? diag : 0.0);
    }
  };
  blk(KL.Q, n, n, v.Q, preg);            // knot.Q = Lxx; Q.diagonal() += preg     (:829, :835)
  blk(KL.S, n, m, v.S, 0.0);
  blk(KL.R, m, m, v.R, preg);            // knot.R = Luu; R.diagonal() += preg     (:831, :836)
  blk(KL.q, n, 1, v.q, 0.0);
  blk(KL.r, m, 1, v.r, 0.0);
  blk(KL.A, n, n, v.A, 0.0);             // knot.A = dd.Jx_                        (:824)
  blk(KL.B, n, m, v.B, 0.0);             // knot.B = dd.Ju_                        (:825)
  if (v.E.ptr) blk(KL.E, n, n, v.E, 0.0);   // knot.E = dd.Jy_                     (:826)
  else                                      // no Jy array: explicit dynamics, E = -I (explicit-dynamics.hxx:32)
    for (int e = tid; e < n * n; e += NT) rec[KL.E + e] = (e % n) == (e / n) ? -1.0 : 0.0;
  blk(KL.f, n, 1, v.f, 0.0);
  blk(KL.C, c, n, v.C, 0.0);
  blk(KL.D, c, m, v.D, 0.0);
  blk(KL.d, c, 1, v.d, 0.0);
  if (h > 0) {
    blk(KL.Gth, h, h, v.Gth, 0.0);
    blk(KL.Gx, n, h, v.Gx, 0.0);
    blk(KL.Gu, m, h, v.Gu, 0.0);
    blk(KL.gamma, h, 1, v.gamma, 0.0);
  }
  if (t == 0) {                           // prob.G0 = id.Jx_; prob.g0 = Lds[0]     (:867-868)
    const int c0 = p.nc0;
    double *G0 = p.G0 + (size_t)b * c0 * n, *g0 = p.g0 + (size_t)b * c0;
    for (int e = tid; e < c0 * n; e += NT) G0[e] = fld(v.G0, b, 0, e % (c0 > 0 ? c0 : 1), e / (c0 > 0 ? c0 : 1));
    for (int i = tid; i < c0; i += NT) g0[i] = fld(v.g0, b, 0, i, 0);
  }
}

int launch_pack_from_device(const DevParams &p, const garb200_device_problem_view &v, double preg, cudaStream_t st) {
  pack_from_device_kernel<<<(unsigned)((size_t)p.batch * (p.T + 1)), 128, 0, st>>>(p, v, preg);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

// ---- trial = cur + alpha * dir over the four trajectory arrays in one launch ---------------
struct Axpy4 { const double *cur[4]; const double *dir[4]; double *out[4]; long long len[4]; };

__global__ void __launch_bounds__(256) linear_step_kernel(const Axpy4 a, double alpha) {
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const double *c = a.cur[k], *d = a.dir[k];
    double *o = a.out[k];
    const long long n = a.len[k];
    if (!o || n == 0) continue;
    if (((reinterpret_cast<uintptr_t>(c) | reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(o)) & 15) == 0) {
      const long long n2 = n >> 1;
      for (long long i = t0; i < n2; i += stride) {
        const double2 x = reinterpret_cast<const double2 *>(c)[i], y = reinterpret_cast<const double2 *>(d)[i];
        reinterpret_cast<double2 *>(o)[i] = make_double2(fma(alpha, y.x, x.x), fma(alpha, y.y, x.y));
      }
      if ((n & 1) && t0 == 0) o[n - 1] = fma(alpha, d[n - 1], c[n - 1]);
    } else {
      for (long long i = t0; i < n; i += stride) o[i] = fma(alpha, d[i], c[i]);
    }
  }
}

int launch_linear_step(const DevParams &p, double alpha, const double *const cur[4], double *const out[4],
                       cudaStream_t st) {
  const long long B = p.batch, T = p.T;
  Axpy4 a;
  const double *dir[4] = {p.xs, p.us, p.vs, p.lbdas};
  const long long len[4] = {B * (T + 1) * p.nx, B * (T + 1) * p.nu, B * (T + 1) * p.nc, B * (p.nc0 + T * p.nx)};
  for (int k = 0; k < 4; k++) { a.cur[k] = cur[k]; a.dir[k] = dir[k]; a.out[k] = (cur[k] && out[k]) ? out[k] : nullptr; a.len[k] = len[k]; }
  linear_step_kernel<<<148 * 8, 256, 0, st>>>(a, alpha);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

// ---- gains-based rollout step ---------------------------------------------------------------
// One CTA per problem streams [ff | fb rows K, Z, L] of every gains record (the A rows and fth are
// never read: (nu + nc + nx) / (nu + nc + 2 nx) of the fb bytes) through an NSTAGE-deep TMA ring;
// 4 threads per row.  dxs is GIVEN for every knot (the caller's nonlinear rollout produces it), so
// knots are independent; one CTA walks its problem's knots for the sake of the ring.
template <int NT, int NSTAGE>
__global__ void __launch_bounds__(NT) rollout_kernel(const DevParams p, double alpha, const double *__restrict__ dxs,
                                                     double *dus, double *dvs, double *dlams) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.nx, T = p.T, c0 = p.nc0, tid = threadIdx.x, b = blockIdx.x;
  const int part = p.GL.fb + (p.nu + p.nc + n) * n;          // doubles of a record that are read (even: n even or ...)
  const int pstride = (part + 1) & ~1;
  double *ring = sm;
  double *x = ring + (size_t)NSTAGE * pstride;
  uint64_t *bars = reinterpret_cast<uint64_t *>(x + ((n + 1) & ~1));
  const size_t xoff = (size_t)b * (T + 1) * n, uoff = (size_t)b * (T + 1) * p.nu, voff = (size_t)b * (T + 1) * p.nc,
               loff = (size_t)b * (c0 + (size_t)T * n);
  const size_t gbase = (size_t)b * (T + 1);
  const uint32_t gbytes = (uint32_t)(pstride * 8);             // even-rounded: still inside the record (the A rows follow)
  if (tid == 0) {
    for (int s = 0; s < NSTAGE; s++) mbar_init(bars + s, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0)
    for (int s = 0; s < NSTAGE && s <= T; s++) {
      mbar_expect_tx(bars + s, gbytes);
      tma_load_1d(ring + (size_t)s * pstride, p.gains + (gbase + phys_t(p, s)) * p.GL.stride, gbytes, bars + s);
    }
  const int q = tid & 3, rl = tid >> 2;
  for (int t = 0; t <= T; t++) {
    const int s = t % NSTAGE;
    for (int i = tid; i < n; i += NT) x[i] = dxs[xoff + (size_t)t * n + i];
    mbar_wait(bars + s, (uint32_t)((t / NSTAGE) & 1));
    __syncthreads();
    const double *rec = ring + (size_t)s * pstride;
    const double *ff = rec + p.GL.ff, *fb = rec + p.GL.fb;
    const int m = p.nu_t[t], c = p.nc_t[t];
    const int nr = (t == T) ? m + c : m + c + n;              // the terminal knot has no L rows (riccati-kernel.hxx:90-150)
    for (int r0 = 0; r0 < nr; r0 += NT / 4) {
      const int r = r0 + rl;
      double acc = 0.0;
      if (r < nr) {
        const double *row = fb + (size_t)r * n;
        for (int j = q; j < n; j += 4) acc = fma(row[j], x[j], acc);
      }
      acc += __shfl_xor_sync(0xffffffffu, acc, 1);
      acc += __shfl_xor_sync(0xffffffffu, acc, 2);
      if (r < nr && q == 0) {
        const double v = fma(alpha, ff[r], acc);
        if (r < m) dus[uoff + (size_t)t * p.nu + r] = v;
        else if (r < m + c) dvs[voff + (size_t)t * p.nc + (r - m)] = v;
        else dlams[loff + c0 + (size_t)t * n + (r - m - c)] = v;     // dlams[t + 1]
      }
    }
    for (int i = m + tid; i < p.nu; i += NT) dus[uoff + (size_t)t * p.nu + i] = 0.0;
    for (int i = c + tid; i < p.nc; i += NT) dvs[voff + (size_t)t * p.nc + i] = 0.0;
    __syncthreads();                                          // all reads of x and of ring slot s done
    if (tid == 0 && t + NSTAGE <= T) {
      mbar_expect_tx(bars + s, gbytes);
      tma_load_1d(ring + (size_t)s * pstride, p.gains + (gbase + phys_t(p, t + NSTAGE)) * p.GL.stride, gbytes, bars + s);
    }
  }
}

int launch_rollout(const DevParams &p, double alpha, const double *dxs, double *dus, double *dvs, double *dlams,
                   cudaStream_t st) {
  constexpr int NT = 128, NSTAGE = 2;
  const int part = p.GL.fb + (p.nu + p.nc + p.nx) * p.nx;
  const size_t smem = ((size_t)NSTAGE * ((part + 1) & ~1) + ((p.nx + 1) & ~1) + NSTAGE + 2) * sizeof(double);
  if (smem > 227u * 1024u) return -1;
  auto kern = rollout_kernel<NT, NSTAGE>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -2;
  kern<<<p.batch, NT, smem, st>>>(p, alpha, dxs, dus, dvs, dlams);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

} // namespace garb200
