"""Synthetic random LQ problems (host side, numpy), packed for the C-ABI.

Distributions restate the reference's test/bench generator
(/root/reference/tests/gar/test_util.cpp:8-66, used by bench/gar-riccati.cpp:23-24):
``[Q S; S^T R] = W W^T`` with ``W ~ N(0,1)^{(nx+nu) x (nx+nu+1)}`` (Wishart), the top-left
``floor(nx/2)^2`` block of ``Q`` zeroed on non-terminal knots (``singular=true``),
``q, r, A ~ N(0,1)``, ``B ~ U[-1,1]``, ``E ~ 1000 N(0,1)``, ``f ~ N(0, 100^2)``,
parameter block ``Gx, Gu, gamma ~ N(0,1)``, ``Gth`` Wishart with ``nth+2`` dof; terminal knot
``nu = 0`` and non-singular; ``G0 = I``, ``g0 = -x0``.

Differences, on purpose: (i) numpy's PCG64 instead of libstdc++'s mt19937 (parity is
defined on identical inputs, the oracle is fed these very bytes); (ii) every (problem,
knot) gets its own draw by default, whereas the reference repeats one knot for all t
(test_util.cpp:56-59) — ``same_knot=True`` restores that; (iii) the reference never sets
``nc > 0``; here ``C, D, d ~ N(0,1)``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .layout import KnotLayout


@dataclass
class PackedProblems:
    """A batch of LQ problems in the packed device format."""
    layout: KnotLayout
    horizon: int            # T  (T+1 knots)
    nc0: int
    nu_t: np.ndarray        # int32 [T+1] per-knot nu
    nc_t: np.ndarray        # int32 [T+1] per-knot nc
    knots: np.ndarray       # float64 [B, T+1, stride]
    G0: np.ndarray          # float64 [B, nc0*nx]  (column-major nc0 x nx)
    g0: np.ndarray          # float64 [B, nc0]

    @property
    def batch(self) -> int:
        return self.knots.shape[0]

    def problem(self, b: int) -> "PackedProblems":
        return PackedProblems(self.layout, self.horizon, self.nc0, self.nu_t, self.nc_t,
                              self.knots[b:b + 1], self.G0[b:b + 1], self.g0[b:b + 1])


def _wishart(rng, lead, n, p):
    w = rng.standard_normal(lead + (n, p))
    return w @ np.swapaxes(w, -1, -2)


def _colmajor(a):
    """[..., r, c] -> [..., r*c] column-major."""
    return np.swapaxes(a, -1, -2).reshape(a.shape[:-2] + (-1,))


def generate_problems(batch, horizon, nx, nu, nc=0, nth=0, *, seed=42, dist="ref",
                      same_knot=False, term_nu=0, term_nc=0, x0=None) -> PackedProblems:
    rng = np.random.default_rng(seed)
    L = KnotLayout(nx, nu, nc, nth)
    off = L.offsets
    T = horizon
    B = batch
    knots = np.zeros((B, T + 1, L.stride))
    nu_t = np.full(T + 1, nu, dtype=np.int32)
    nc_t = np.full(T + 1, nc, dtype=np.int32)
    nu_t[T] = term_nu
    nc_t[T] = term_nc

    def put(sl, name, arr):
        r, c = arr.shape[-2:]
        if r * c:
            knots[:, sl, off[name]: off[name] + r * c] = _colmajor(arr)

    def gen(nt, m, c, singular):
        lead = (B, nt)
        f = {}
        qsr = _wishart(rng, lead, nx + m, nx + m + 1)
        Q = qsr[..., :nx, :nx].copy()
        if singular and dist != "explicit_psd":   # "explicit_psd": convex stage costs (value function stays PSD)
            h = nx // 2
            Q[..., :h, :h] = 0.0
        f["Q"], f["S"], f["R"] = Q, qsr[..., :nx, nx:], qsr[..., nx:, nx:]
        f["q"] = rng.standard_normal(lead + (nx, 1))
        f["r"] = rng.standard_normal(lead + (m, 1))
        if dist == "ref":
            f["A"] = rng.standard_normal(lead + (nx, nx))
            f["E"] = 1000.0 * rng.standard_normal(lead + (nx, nx))
        elif dist in ("explicit", "explicit_psd"):
            # what ProxDDP feeds: explicit dynamics, E = -I (explicit-dynamics.hxx:32)
            f["A"] = 0.9 * np.eye(nx) + (0.1 / np.sqrt(nx)) * rng.standard_normal(lead + (nx, nx))
            f["E"] = np.broadcast_to(-np.eye(nx), lead + (nx, nx)).copy()
        else:
            raise ValueError(f"unknown dist {dist!r}")
        f["B"] = rng.uniform(-1.0, 1.0, lead + (nx, m))
        f["f"] = 100.0 * rng.standard_normal(lead + (nx, 1))
        f["C"] = rng.standard_normal(lead + (c, nx))
        f["D"] = rng.standard_normal(lead + (c, m))
        f["d"] = rng.standard_normal(lead + (c, 1))
        if nth > 0:
            f["Gx"] = rng.standard_normal(lead + (nx, nth))
            f["Gu"] = rng.standard_normal(lead + (m, nth))
            f["Gth"] = _wishart(rng, lead, nth, nth + 2)
            f["gamma"] = rng.standard_normal(lead + (nth, 1))
        return f

    if T > 0:
        nt = 1 if same_knot else T
        f = gen(nt, nu, nc, True)
        for name, arr in f.items():
            if same_knot:
                arr = np.broadcast_to(arr, (B, T) + arr.shape[2:])
            put(slice(0, T), name, arr)
    ft = gen(1, term_nu, term_nc, False)
    for name, arr in ft.items():
        put(slice(T, T + 1), name, arr)

    nc0 = nx
    if x0 is None:
        x0 = rng.standard_normal((B, nx))
    x0 = np.broadcast_to(np.asarray(x0, dtype=np.float64), (B, nx))
    G0 = np.broadcast_to(np.eye(nx).reshape(-1), (B, nx * nx)).copy()
    g0 = -x0.copy()
    return PackedProblems(L, T, nc0, nu_t, nc_t, knots, G0, g0)
