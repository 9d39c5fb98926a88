"""CPU tests of the oracle (oracle/): it must satisfy the reference's own assertions and
agree with an independent dense LAPACK solve before any GPU number is trusted.

Mirrors /root/reference/tests/gar/riccati.cpp, parallel.cpp, block-matrix.cpp.
"""
import numpy as np
import pytest
import scipy.linalg as sl

from aligator_swk_b200.layout import GainLayout, KnotLayout, ValueLayout
from aligator_swk_b200.synth import generate_problems
from conftest import relerr


def dense_solve(oracle, pp, mu, b=0):
    mat, rhs = oracle.dense_kkt(pp, mu, mu, b)
    z = sl.solve(mat, -rhs, assume_a="sym")
    return oracle.dense_solution_to_traj(pp, z)


def test_layout_mirror_matches_c_header(oracle):
    for dims in [(36, 12, 0, 0), (12, 4, 6, 4), (3, 2, 1, 1), (76, 32, 0, 0), (5, 0, 3, 2)]:
        k, g, v = oracle.layouts(*dims)
        o = KnotLayout(*dims).offsets
        names = ("Q", "S", "R", "q", "r", "A", "B", "E", "f", "C", "D", "d", "Gth", "Gx", "Gu", "gamma", "stride")
        assert [o[n] for n in names] == list(k)
        G = GainLayout(*dims)
        assert [G.nrows, G.offsets["ff"], G.offsets["fb"], G.offsets["fth"], G.stride] == list(g)
        V = ValueLayout(dims[0], dims[3]).offsets
        assert [V[n] for n in ("Pmat", "pvec", "Vxx", "vx", "Vxt", "Vtt", "vt", "stride")] == list(v)


@pytest.mark.parametrize("n", [1, 2, 5, 12, 33, 72])
def test_factorizations_vs_lapack(oracle, n):
    import ctypes as C
    rng = np.random.default_rng(n)
    L = oracle.lib()
    dp = C.POINTER(C.c_double)
    A = rng.standard_normal((n, n))
    X = rng.standard_normal((n, 3))
    # LU
    Xc = np.asfortranarray(X.copy())
    assert L.gar_oracle_lu_solve(n, 3, np.asfortranarray(A).ctypes.data_as(dp), Xc.ctypes.data_as(dp)) == 0
    assert relerr(Xc, np.linalg.solve(A, X)) < 1e-9
    # LLT
    S = A @ A.T + n * np.eye(n)
    Xc = np.asfortranarray(X.copy())
    assert L.gar_oracle_llt_solve(n, 3, np.asfortranarray(S).ctypes.data_as(dp), Xc.ctypes.data_as(dp)) == 0
    assert relerr(Xc, np.linalg.solve(S, X)) < 1e-11
    # Bunch-Kaufman on a symmetric INDEFINITE saddle matrix (forces 2x2 pivots)
    m = n // 2
    K = np.zeros((n, n))
    H = rng.standard_normal((n - m, n - m)); K[: n - m, : n - m] = H @ H.T * 1e-3
    J = rng.standard_normal((m, n - m)); K[n - m:, : n - m] = J; K[: n - m, n - m:] = J.T
    K[n - m:, n - m:] = -1e-8 * np.eye(m)
    K += 1e-6 * np.eye(n) if m == 0 else 0
    Xc = np.asfortranarray(X.copy())
    piv = np.zeros(n, np.int32)
    assert L.gar_oracle_bk_solve(n, 3, np.asfortranarray(K).ctypes.data_as(dp), Xc.ctypes.data_as(dp),
                                 piv.ctypes.data_as(C.POINTER(C.c_int))) == 0
    ref = np.linalg.solve(K, X)
    assert relerr(Xc, ref) < 1e-7 * max(1.0, np.linalg.cond(K) * 1e-10)
    if n >= 5:
        assert (piv < 0).any(), "expected at least one 2x2 pivot on a saddle matrix"


def test_block_tridiag_vs_dense(oracle):
    """tests/gar/block-matrix.cpp:54-112 — N=6 super-diagonal blocks of size 2, isApprox 1e-12."""
    import ctypes as C
    rng = np.random.default_rng(0)
    nb, bs = 7, 2
    Bm = rng.standard_normal((bs, bs))
    diag = []
    for _ in range(nb):
        w = rng.standard_normal((bs, bs + 1)); diag.append(w @ w.T)
    dense = np.zeros((nb * bs, nb * bs))
    for i in range(nb):
        dense[i * bs:(i + 1) * bs, i * bs:(i + 1) * bs] = diag[i]
        if i + 1 < nb:
            dense[i * bs:(i + 1) * bs, (i + 1) * bs:(i + 2) * bs] = Bm
            dense[(i + 1) * bs:(i + 2) * bs, i * bs:(i + 1) * bs] = Bm.T
    rhs = np.ones(nb * bs)
    d = np.concatenate([m.reshape(-1, order="F") for m in diag])
    s = np.concatenate([Bm.reshape(-1, order="F")] * (nb - 1))
    dp = C.POINTER(C.c_double)
    x = rhs.copy()
    assert oracle.lib().gar_oracle_blocktridiag_solve(nb, bs, d.ctypes.data_as(dp), s.ctypes.data_as(dp), x.ctypes.data_as(dp)) == 0
    assert np.allclose(x, np.linalg.solve(dense, rhs), rtol=1e-12, atol=0)


def test_one_knot_prob(oracle):
    """tests/gar/riccati.cpp:93-111"""
    pp = generate_problems(1, 0, 2, 2, seed=3, x0=np.zeros(2))
    out = oracle.prox_solve(pp, 1e-13, 1e-13)
    assert oracle.kkt_error(pp, out["xs"], out["us"], out["vs"], out["lbdas"], 0.0, 0.0).max() <= 1e-10


def test_random_long_problem(oracle):
    """tests/gar/riccati.cpp:113-158: nx=36 nu=12 N=100, same knot repeated, mu=1e-14, KKT<=1e-9"""
    pp = generate_problems(1, 100, 36, 12, seed=42, same_knot=True, x0=np.zeros(36))
    out = oracle.prox_solve(pp, 1e-14, 1e-14)
    assert oracle.kkt_error(pp, out["xs"], out["us"], out["vs"], out["lbdas"], 0.0, 0.0).max() <= 1e-9
    xs, us, vs, lb = dense_solve(oracle, pp, 1e-14)
    assert relerr(out["xs"], xs) < 1e-9 and relerr(out["us"], us) < 1e-9 and relerr(out["lbdas"], lb) < 1e-9


def test_parametric(oracle):
    """tests/gar/riccati.cpp:160-205: nx=3 nu=2 N=100 nth=1 mu=1e-12, KKT <= 1e-10 with theta"""
    rng = np.random.default_rng(5)
    pp = generate_problems(1, 100, 3, 2, nth=1, seed=7, same_knot=True)
    theta = rng.uniform(-1, 1, 1)
    out = oracle.prox_solve(pp, 1e-12, 1e-12, theta=theta)
    e = oracle.kkt_error(pp, out["xs"], out["us"], out["vs"], out["lbdas"], 0.0, 0.0, theta=theta)
    assert e.max() <= 1e-10
    for k in ("kkt0_ff", "kkt0_fth", "thGrad", "thHess"):
        assert np.isfinite(out[k]).all()


def short_horz_pb():
    """tests/gar/riccati.cpp:20-58 (B, f random there too)."""
    from aligator_swk_b200 import gar
    rng = np.random.default_rng(11)
    nx = nu = 2
    x0, x1 = np.ones(nx), -np.ones(nx)

    def init_knot(nc=0):
        k = gar.LqrKnot(nx, nu, nc)
        k.A[:] = [[0.1, 0.0], [-0.1, 0.01]]
        k.B[:] = rng.uniform(-1, 1, (nx, nu))
        k.E[:] = -np.eye(nx)
        k.f[:] = rng.uniform(-1, 1, nx)
        k.Q[:] = 0.01 * np.eye(nx)
        k.R[:] = 0.1 * np.eye(nu)
        return k
    base = init_knot()
    knots = [base.copy() for _ in range(9)]
    knots[4] = init_knot(nu); knots[4].D[:] = np.eye(nu); knots[4].d[:] = 0.1
    knots[8] = base.copy(); knots[8].Q[:] = np.eye(nx); knots[8].q[:] = -x1
    prob = gar.LqrProblem(knots, nx)
    prob.g0[:] = -x0
    prob.G0[:] = np.eye(nx)
    return prob


def test_short_horz_pb(oracle):
    """ragged nc (one knot with nc = nu), E = -I, KKT <= 1e-9"""
    pp = short_horz_pb().pack()
    out = oracle.prox_solve(pp, 1e-14, 1e-14)
    assert oracle.kkt_error(pp, out["xs"], out["us"], out["vs"], out["lbdas"], 0.0, 0.0).max() <= 1e-9
    xs, us, vs, lb = dense_solve(oracle, pp, 1e-14)
    assert relerr(out["xs"], xs) < 1e-9 and relerr(out["us"], us) < 1e-9


@pytest.mark.parametrize("dims", [(12, 4, 0, 50), (36, 12, 6, 30), (7, 3, 2, 9), (5, 5, 5, 12)])
def test_prox_vs_dense_lapack(oracle, dims):
    nx, nu, nc, T = dims
    pp = generate_problems(2, T, nx, nu, nc, seed=nx + T)
    for b in range(2):
        out = oracle.prox_solve(pp, 1e-10, 1e-10, b=b)
        xs, us, vs, lb = dense_solve(oracle, pp, 1e-10, b)
        for k, ref in (("xs", xs), ("us", us), ("vs", vs), ("lbdas", lb)):
            assert relerr(out[k], ref) < 1e-9, (dims, k)


def test_parallel_solver_class(oracle):
    """tests/gar/parallel.cpp:199-271: nx=32 nu=12 N=50, 6 legs, mu=1e-9; serial == parallel 1e-10"""
    pp = generate_problems(1, 50, 32, 12, seed=42, same_knot=True, x0=np.zeros(32))
    mu, tol = 1e-9, 1e-10
    ref = oracle.prox_solve(pp, mu, mu)
    assert oracle.kkt_error(pp, ref["xs"], ref["us"], ref["vs"], ref["lbdas"], mu, mu).max() <= tol
    par = oracle.parallel_solve(pp, mu, mu, 6)
    assert par["info"][1] == 1
    assert par["kkt"].max() <= tol
    assert np.abs(par["xs"] - ref["xs"]).max() <= tol
    assert np.abs(par["lbdas"] - ref["lbdas"]).max() <= tol


@pytest.mark.parametrize("legs", [2, 3, 4, 16])
def test_parallel_legs_sweep(oracle, legs):
    pp = generate_problems(1, 40, 6, 3, seed=legs)
    mu = 1e-9
    ref = oracle.prox_solve(pp, mu, mu)
    par = oracle.parallel_solve(pp, mu, mu, legs)
    assert relerr(par["xs"], ref["xs"]) < 1e-9 and relerr(par["us"], ref["us"]) < 1e-9
    assert relerr(par["lbdas"], ref["lbdas"]) < 1e-9


def test_golden_vectors(oracle):
    """Committed fixtures (tests/golden/make_golden.py): oracle today == oracle when committed,
    and both == the dense LAPACK solution stored beside them."""
    import os
    gd = os.path.join(os.path.dirname(__file__), "golden")
    files = sorted(f for f in os.listdir(gd) if f.endswith(".npz"))
    assert files, "no golden fixtures"
    from golden.make_golden import load_case
    for f in files:
        pp, mu, theta, gold = load_case(os.path.join(gd, f))
        out = oracle.prox_solve(pp, mu, mu, theta=theta)
        for k in ("xs", "us", "vs", "lbdas"):
            assert relerr(out[k], gold[k]) < 1e-12, (f, k)
            if theta is None:
                assert relerr(out[k], gold["dense_" + k]) < 1e-9, (f, k)


@pytest.mark.parametrize("dims", [(36, 12, 0, 0, 100, 1e-14), (12, 4, 3, 2, 30, 1e-10), (3, 2, 0, 1, 100, 1e-12)],
                         ids=lambda d: "nx%d_nu%d_nc%d_nth%d_T%d_mu%g" % d)
def test_dense_solver_restatement(oracle, dims):
    """RiccatiSolverDense (dense-kernel.hpp:56-212, dense-riccati.hxx:46-118) restated in oracle/dense.cpp: the
    reference's own check (tests/gar/riccati.cpp:143-157) — KKT <= 1e-9 and the same solution as the proximal
    solver — plus agreement with the dense LAPACK solve of the lqrDenseMatrix KKT system."""
    nx, nu, nc, nth, T, mu = dims
    pp = generate_problems(2, T, nx, nu, nc, nth, seed=nx + T, same_knot=(nx == 36), **({"x0": np.zeros(nx)} if nx == 36 else {}))
    th = np.random.default_rng(1).standard_normal(nth) if nth else None
    for b in range(2):
        d = oracle.dense_solve(pp, mu, mu, theta=th, b=b)
        a = oracle.prox_solve(pp, mu, mu, theta=th, b=b)
        assert d["kkt"].max() <= 1e-9
        for k in ("xs", "us", "vs", "lbdas"):
            assert relerr(d[k], a[k]) < 1e-9, (dims, b, k)
        if nth == 0:
            xs, us, vs, lb = dense_solve(oracle, pp, mu, b)
            assert relerr(d["xs"], xs) < 1e-9 and relerr(d["us"], us) < 1e-9


@pytest.mark.parametrize("dims", [(5, 2, 1, 1, 7, 12, 4, 3, 2), (7, 3, 0, 0, 9, 12, 4, 0, 0), (6, 2, 2, 0, 5, 8, 5, 4, 1)],
                         ids=lambda d: "nx%d_nu%d_nc%d_nth%d_T%d_into_%d_%d_%d_%d" % d)
def test_padding_embedding_is_exact(oracle, dims):
    """The embedding csrc/pad.cu uses for shapes without an instantiated sweep, checked on the CPU
    independently of any kernel: pad a problem with E' = [E 0; 0 -I], no cost on the padded states, R' = [R 0; 0 I], empty
    constraint rows, zero theta columns, G0' = [G0 0; 0 I]; the oracle's solution of the padded problem restricted
    to the original rows IS the oracle's solution of the original problem (to round-off), the padded parts are 0."""
    nx, nu, nc, nth, T, NX, NU, NC, NTH = dims
    mu = 1e-8
    pp = generate_problems(1, T, nx, nu, nc, nth, seed=17)
    L, LP = pp.layout, KnotLayout(NX, NU, NC, NTH)
    knots = np.zeros((1, T + 1, LP.stride))
    for t in range(T + 1):
        m, c = int(pp.nu_t[t]), int(pp.nc_t[t])
        M, Cc = (NU, NC) if t < T else (0, 0)
        f = L.unpack(pp.knots[0, t], m, c)
        g = {k: np.zeros(s) for k, s in LP.shapes(M, Cc).items()}
        for k, a in f.items():
            g[k][:a.shape[0], :a.shape[1]] = a
        for i in range(nx, NX):
            g["E"][i, i] = -1.0
        for i in range(m, M):
            g["R"][i, i] = 1.0
        LP.pack(knots[0, t], g, M, Cc)
    c0 = pp.nc0
    C0 = c0 + NX - nx
    G0 = np.zeros((C0, NX)); G0[:c0, :nx] = pp.G0[0].reshape(nx, c0).T
    G0[c0:, nx:] = np.eye(NX - nx)
    g0 = np.concatenate([pp.g0[0], np.zeros(NX - nx)])
    nu_t = np.array([NU] * T + [0], np.int32); nc_t = np.array([NC] * T + [0], np.int32)
    padded = type(pp)(LP, T, C0, nu_t, nc_t, knots, np.ascontiguousarray(G0.T).reshape(1, -1), g0.reshape(1, -1))
    theta = np.random.default_rng(2).standard_normal(nth) if nth else None
    thp = None if NTH == 0 else np.concatenate([theta if theta is not None else np.zeros(0), np.zeros(NTH - nth)])
    a = oracle.prox_solve(pp, mu, mu, theta=theta)
    b = oracle.prox_solve(padded, mu, mu, theta=thp)
    xs, us, vs = b["xs"], b["us"], b["vs"]
    assert relerr(xs[:, :nx], a["xs"]) < 1e-11 and np.abs(xs[:, nx:]).max(initial=0.0) < 1e-12
    if nu:
        assert relerr(us[:, :nu], a["us"]) < 1e-11
    assert np.abs(us[:, nu:]).max(initial=0.0) < 1e-12
    if nc:
        assert relerr(vs[:, :nc], a["vs"]) < 1e-10
    lb = b["lbdas"]
    assert relerr(lb[:c0], a["lbdas"][:c0]) < 1e-10
    assert relerr(lb[C0:].reshape(T, NX)[:, :nx], a["lbdas"][c0:].reshape(T, nx)) < 1e-10
    assert np.abs(lb[C0:].reshape(T, NX)[:, nx:]).max(initial=0.0) < 1e-12
