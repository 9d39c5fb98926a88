// pad.cu — shape generality by EXACT padding: a problem whose (nx, nu, nc, nth) has no
// instantiated sweep is embedded in the next instantiated shape (NX, NU, NC, NTH) >= (nx, nu, nc, nth):
//   padded states    E' = [E 0; 0 -I], A' = [A 0; 0 0], B' = [B 0; 0 0], f' = [f; 0], Q' = [Q 0; 0 0]:
//                    x_pad(t+1) = 0 by their dynamics, no cost -> they stay at zero and decouple exactly
//   padded controls  R' = [R 0; 0 I], S', r' zero           -> u_pad = 0
//   padded rows of C x + D u + d = 0 are empty               -> pivot -mueq, multiplier 0
//   padded theta     all blocks zero
//   initial stage    G0' = [G0 0; 0 I], g0' = [g0; 0]        -> x_pad(0) = 0 (G0 == I stays == I)
// Every cross term between the real and the padded part is an exact zero, so the real part of the padded
// solve is the solve of the original problem (up to the different grouping of exact-zero terms).
// The caller-visible layouts stay the USER's: these kernels convert between them and the internal ones.
#include "kernels.cuh"

namespace garb200 {

// u: user-shaped params (problem in user layout), p: internal padded params
__global__ void __launch_bounds__(128) pad_problem_kernel(const DevParams u, const DevParams p) {
  constexpr int NT = 128;
  const int T = p.T, tid = threadIdx.x;
  const int b = blockIdx.x / (T + 1), t = blockIdx.x % (T + 1);
  const int n = u.nx, N = p.nx, h = u.nth, H = p.nth;
  const int m = u.nu_t[t], c = u.nc_t[t];          // this knot's own dims (user)
  const int M = p.nu_t[t], C_ = p.nc_t[t];         // ... internal (0 at the terminal knot, else NU / NC)
  const garb200_knot_layout &UL = u.KL, &PL = p.KL;
  const double *src = u.knots + ((size_t)b * (T + 1) + phys_t(u, t)) * UL.stride;
  double *dst = p.knots + ((size_t)b * (T + 1) + phys_t(p, t)) * PL.stride;
  for (int e = tid; e < PL.stride; e += NT) dst[e] = 0.0;
  __syncthreads();
  // copy an (r x cc) block with leading dims (lds -> ldd); rows/cols beyond it stay zero
  auto blk = [&](int so, int lds, int dof, int ldd, int r, int cc) {
    for (int e = tid; e < r * cc; e += NT) { const int i = e % r, j = e / r; dst[dof + i + j * ldd] = src[so + i + j * lds]; }
  };
  blk(UL.Q, n, PL.Q, N, n, n);
  // (Q_pad stays 0: the padded states are pinned by their constraints alone -- x_pad(0) = 0 through G0', x_pad(t+1) = 0 through
  //  E_pad = -I with A_pad = B_pad = 0 --, and a unit cost block would dominate ||Ptilde||_F, which selects the order of the
  //  Neumann series for Lambda^{-1}: measured +10 % (36,12) / +40 % (76,32) on the padded shapes with Q_pad = I)
  blk(UL.S, n, PL.S, N, n, m);
  blk(UL.R, m, PL.R, M, m, m);
  for (int i = m + tid; i < M; i += NT) dst[PL.R + i + i * M] = 1.0;           // R_pad = I
  blk(UL.q, n, PL.q, N, n, 1);
  blk(UL.r, m, PL.r, M, m, 1);
  blk(UL.A, n, PL.A, N, n, n);
  blk(UL.B, n, PL.B, N, n, m);
  blk(UL.E, n, PL.E, N, n, n);
  for (int i = n + tid; i < N; i += NT) dst[PL.E + i + i * N] = -1.0;          // E_pad = -I
  blk(UL.f, n, PL.f, N, n, 1);
  blk(UL.C, c, PL.C, C_, c, n);
  blk(UL.D, c, PL.D, C_, c, m);
  blk(UL.d, c, PL.d, C_, c, 1);
  if (h > 0) {
    blk(UL.Gth, h, PL.Gth, H, h, h);
    blk(UL.Gx, n, PL.Gx, N, n, h);
    blk(UL.Gu, m, PL.Gu, M, m, h);
    blk(UL.gamma, h, PL.gamma, H, h, 1);
  }
  if (t == 0) {
    const int c0 = u.nc0, C0 = p.nc0;
    const double *G0 = u.G0 + (size_t)b * c0 * n, *g0 = u.g0 + (size_t)b * c0;
    double *G0p = p.G0 + (size_t)b * C0 * N, *g0p = p.g0 + (size_t)b * C0;
    for (int e = tid; e < C0 * N; e += NT) {
      const int i = e % C0, j = e / C0;
      double v = 0.0;
      if (i < c0 && j < n) v = G0[i + j * c0];
      else if (i >= c0 && j >= n && (i - c0) == (j - n)) v = 1.0;
      G0p[e] = v;
    }
    for (int i = tid; i < C0; i += NT) g0p[i] = i < c0 ? g0[i] : 0.0;
  }
}

// internal gains records -> user gains records (block rows [K; Z; L; A], row-major fb / fth)
__global__ void __launch_bounds__(128) unpad_gains_kernel(const DevParams u, const DevParams p) {
  constexpr int NT = 128;
  const int T = p.T, tid = threadIdx.x;
  const int b = blockIdx.x / (T + 1), t = blockIdx.x % (T + 1);
  const int n = u.nx, N = p.nx, hu = u.GL.nth, hp = p.GL.nth;
  const int m = u.nu_t[t], c = u.nc_t[t], M = p.nu_t[t], C_ = p.nc_t[t];
  const double *src = p.gains + ((size_t)b * (T + 1) + phys_t(p, t)) * p.GL.stride;
  double *dst = u.gains + ((size_t)b * (T + 1) + phys_t(u, t)) * u.GL.stride;
  const bool term_empty = (t == T);               // the terminal knot has block rows [K; Z] only
  const int nr = term_empty ? m + c : m + c + 2 * n;
  auto prow = [&](int r) {                        // user row -> internal row
    if (r < m) return r;
    if (r < m + c) return M + (r - m);
    if (r < m + c + n) return M + C_ + (r - m - c);
    return M + C_ + N + (r - m - c - n);
  };
  for (int e = tid; e < u.GL.stride; e += NT) {
    double v = 0.0;
    if (e >= u.GL.fth) { if (hu) { const int k = e - u.GL.fth, r = k / hu, a = k % hu; if (r < nr) v = src[p.GL.fth + prow(r) * hp + a]; } }
    else if (e >= u.GL.fb) { const int k = e - u.GL.fb, r = k / n, j = k % n; if (r < nr) v = src[p.GL.fb + prow(r) * N + j]; }
    else if (e < nr) v = src[p.GL.ff + prow(e)];
    dst[e] = v;
  }
}

// internal trajectories -> user trajectories (any output may be null)
__global__ void __launch_bounds__(256) unpad_traj_kernel(const DevParams u, const DevParams p, double *xs, double *us,
                                                         double *vs, double *lbdas) {
  const int T = p.T, b = blockIdx.x, tid = threadIdx.x, NT = 256;
  const int n = u.nx, N = p.nx, m = u.nu, M = p.nu, c = u.nc, C_ = p.nc, c0 = u.nc0, C0 = p.nc0;
  if (xs) for (int e = tid; e < (T + 1) * n; e += NT) xs[(size_t)b * (T + 1) * n + e] = p.xs[(size_t)b * (T + 1) * N + (e / n) * N + e % n];
  if (us && m) for (int e = tid; e < (T + 1) * m; e += NT) us[(size_t)b * (T + 1) * m + e] = p.us[(size_t)b * (T + 1) * M + (e / m) * M + e % m];
  if (vs && c) for (int e = tid; e < (T + 1) * c; e += NT) vs[(size_t)b * (T + 1) * c + e] = p.vs[(size_t)b * (T + 1) * C_ + (e / c) * C_ + e % c];
  if (lbdas) {
    const size_t lo = (size_t)b * (c0 + (size_t)T * n), lp = (size_t)b * (C0 + (size_t)T * N);
    for (int i = tid; i < c0; i += NT) lbdas[lo + i] = p.lbdas[lp + i];
    for (int e = tid; e < T * n; e += NT) lbdas[lo + c0 + e] = p.lbdas[lp + C0 + (size_t)(e / n) * N + e % n];
  }
}

// user theta [batch][nth] -> internal [batch][NTH], zero padded
__global__ void pad_theta_kernel(const double *th, double *thp, int batch, int h, int H) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < batch * H) { const int b = e / H, a = e % H; thp[e] = a < h ? th[b * h + a] : 0.0; }
}

int launch_pad_problem(const DevParams &u, const DevParams &p, cudaStream_t st) {
  pad_problem_kernel<<<(unsigned)((size_t)p.batch * (p.T + 1)), 128, 0, st>>>(u, p);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}
int launch_unpad_gains(const DevParams &u, const DevParams &p, cudaStream_t st) {
  unpad_gains_kernel<<<(unsigned)((size_t)p.batch * (p.T + 1)), 128, 0, st>>>(u, p);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}
int launch_unpad_traj(const DevParams &u, const DevParams &p, double *xs, double *us, double *vs, double *lbdas,
                      cudaStream_t st) {
  unpad_traj_kernel<<<p.batch, 256, 0, st>>>(u, p, xs, us, vs, lbdas);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}
int launch_pad_theta(const double *th, double *thp, int batch, int h, int H, cudaStream_t st) {
  if (batch * H == 0) return 0;
  pad_theta_kernel<<<(batch * H + 255) / 256, 256, 0, st>>>(th, thp, batch, h, H);
  return cudaGetLastError() == cudaSuccess ? 1 : -2;
}

/// Smallest instantiated shape that covers (nx, nu, nc, nth); false if none.
bool padded_shape_for(int nx, int nu, int nc, int nth, int out[4]) {
  static const int shapes[][4] = {{12, 4, 0, 0}, {36, 12, 0, 0}, {76, 32, 0, 0}, {36, 12, 6, 4}, {36, 12, 12, 8}};
  for (const auto &s : shapes)
    if (nx <= s[0] && nu <= s[1] && nc <= s[2] && nth <= s[3]) { for (int i = 0; i < 4; i++) out[i] = s[i]; return true; }
  return false;
}

} // namespace garb200
