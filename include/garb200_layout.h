/* garb200_layout.h — packed record layouts shared by the C-ABI, the CUDA kernels,
 * the host mirror and the test oracle.  Plain C, header-only, no dependencies.
 *
 * The reference keeps every knot matrix in its own allocation
 * (`LqrKnotTpl`, /root/reference/include/aligator/gar/lqr-problem.hpp:45-56, sizes
 * lqr-problem.hxx:38-54) and every factor in its own Eigen object
 * (`StageFactor`, riccati-kernel.hpp:35-91).  On the GPU one knot is ONE contiguous,
 * 16-byte aligned record so that a single 1-D bulk copy (TMA `cp.async.bulk`) stages
 * it into shared memory.
 *
 * All sub-blocks are column-major with leading dimension = their own row count,
 * EXCEPT the gain blocks fb / fth which are row-major like the reference
 * (riccati-kernel.hpp:77-78), so getFeedback() needs no transpose.
 *
 * Offsets are computed from the MAXIMUM stage dims (nx, nu, nc, nth) of the solver so
 * that every knot record has the same stride; a knot whose own nu_t / nc_t is smaller
 * (e.g. the terminal knot with nu = 0, tests/gar/test_util.cpp:60) stores its blocks
 * compactly at the start of each region (leading dimension = its own row count).
 */
#ifndef GARB200_LAYOUT_H
#define GARB200_LAYOUT_H

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__CUDACC__)
#define GARB200_HD __host__ __device__
#else
#define GARB200_HD
#endif

/* Offsets (in doubles) of each field inside one packed knot record. */
typedef struct garb200_knot_layout {
  int nx, nu, nc, nth;
  int Q, S, R, q, r;  /* cost:      Q nx*nx, S nx*nu, R nu*nu, q nx, r nu          */
  int A, B, E, f;     /* dynamics:  E x' + A x + B u + f = 0 (lqr-problem.hpp:26-30) */
  int C, D, d;        /* equality:  C x + D u + d = 0                               */
  int Gth, Gx, Gu, gamma; /* parameter block (lqr-problem.hpp:51-55); Gv is NOT stored:
                             the proximal kernel never reads it (riccati-kernel.hxx:128,272) */
  int stride;         /* record length in doubles, multiple of 2 (16 bytes)         */
} garb200_knot_layout;

/* Offsets (in doubles) inside one gains record = the reference's StageFactor::{ff,fb,fth}
 * (riccati-kernel.hpp:76-78): block rows [K; Z; L(=Xi); A] of sizes [nu; nc; nx; nx]. */
typedef struct garb200_gain_layout {
  int nx, nu, nc, nth;
  int nrows;  /* nu + nc + 2 nx                                  */
  int ff;     /* nrows                                           */
  int fb;     /* nrows x nx,  ROW-major                          */
  int fth;    /* nrows x nth, ROW-major                          */
  int stride; /* multiple of 2                                   */
} garb200_gain_layout;

/* Offsets inside one value-function record = StageFactor::value_t
 * (riccati-kernel.hpp:39-60). */
typedef struct garb200_value_layout {
  int nx, nth;
  int Pmat, pvec, Vxx, vx, Vxt, Vtt, vt;
  int stride;
} garb200_value_layout;

static inline GARB200_HD int garb200_round2(int x) { return (x + 1) & ~1; }

static inline GARB200_HD garb200_knot_layout
garb200_make_knot_layout(int nx, int nu, int nc, int nth) {
  garb200_knot_layout L;
  int o = 0;
  L.nx = nx; L.nu = nu; L.nc = nc; L.nth = nth;
  /* every matrix block starts 16-byte aligned */
  L.Q = o;     o = garb200_round2(o + nx * nx);
  L.S = o;     o = garb200_round2(o + nx * nu);
  L.R = o;     o = garb200_round2(o + nu * nu);
  L.q = o;     o = garb200_round2(o + nx);
  L.r = o;     o = garb200_round2(o + nu);
  L.A = o;     o = garb200_round2(o + nx * nx);
  L.B = o;     o = garb200_round2(o + nx * nu);
  L.E = o;     o = garb200_round2(o + nx * nx);
  L.f = o;     o = garb200_round2(o + nx);
  L.C = o;     o = garb200_round2(o + nc * nx);
  L.D = o;     o = garb200_round2(o + nc * nu);
  L.d = o;     o = garb200_round2(o + nc);
  L.Gth = o;   o = garb200_round2(o + nth * nth);
  L.Gx = o;    o = garb200_round2(o + nx * nth);
  L.Gu = o;    o = garb200_round2(o + nu * nth);
  L.gamma = o; o = garb200_round2(o + nth);
  L.stride = o;
  return L;
}

static inline GARB200_HD garb200_gain_layout
garb200_make_gain_layout(int nx, int nu, int nc, int nth) {
  garb200_gain_layout L;
  int o = 0;
  L.nx = nx; L.nu = nu; L.nc = nc; L.nth = nth;
  L.nrows = nu + nc + 2 * nx;
  L.ff = o;  o = garb200_round2(o + L.nrows);
  L.fb = o;  o = garb200_round2(o + L.nrows * nx);
  L.fth = o; o = garb200_round2(o + L.nrows * nth);
  L.stride = o;
  return L;
}

static inline GARB200_HD garb200_value_layout
garb200_make_value_layout(int nx, int nth) {
  garb200_value_layout L;
  int o = 0;
  L.nx = nx; L.nth = nth;
  L.Pmat = o; o = garb200_round2(o + nx * nx);
  L.pvec = o; o = garb200_round2(o + nx);
  L.Vxx = o;  o = garb200_round2(o + nx * nx);
  L.vx = o;   o = garb200_round2(o + nx);
  L.Vxt = o;  o = garb200_round2(o + nx * nth);
  L.Vtt = o;  o = garb200_round2(o + nth * nth);
  L.vt = o;   o = garb200_round2(o + nth);
  L.stride = o;
  return L;
}

#ifdef __cplusplus
}
#endif
#endif /* GARB200_LAYOUT_H */
