"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same
bytes.  FP64 tolerance from BASELINE.json north_star: relative 1e-10 on (xs, us, vs, lbdas)
and gains, KKT residuals reported/asserted the way the reference's tests do
(/root/reference/tests/gar/riccati.cpp: <= 1e-9 / 1e-10).
"""
import os

import numpy as np
import pytest

from aligator_swk_b200.layout import GainLayout
from aligator_swk_b200.synth import generate_problems
from conftest import relerr

pytestmark = pytest.mark.gpu

RTOL = 1e-10          # north_star: "within a relative 1e-10 in fp64"
RTOL_GAINS = 1e-9     # gains of ill-conditioned knots amplify rounding; still far below test tolerances


@pytest.fixture(scope="module")
def capi():
    from aligator_swk_b200 import capi as c
    c.lib()
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return c


def solve_gpu(capi, pp, mu, theta=None, num_legs=0, flags=0):
    s = capi.Solver.for_problems(pp, num_legs=num_legs, flags=flags)
    s.upload(pp)
    s.backward(mu, mu)
    xs, us, vs, lb = s.forward(theta)
    return s, xs, us, vs, lb


def compare_traj(oracle, pp, mu, xs, us, vs, lb, theta=None, legs=0, rtol=RTOL, problems=None):
    worst = 0.0
    for b in (problems if problems is not None else range(pp.batch)):
        ref = (oracle.parallel_solve(pp, mu, mu, legs, b=b) if legs >= 2
               else oracle.prox_solve(pp, mu, mu, theta=theta, b=b))
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            e = relerr(got, ref[key])
            worst = max(worst, e)
            assert e < rtol, (b, key, e)
    return worst


CASES = [
    # nx, nu, nc, nth, T, batch, mu, dist
    (2, 2, 0, 0, 0, 3, 1e-13, "ref"),        # one_knot_prob (riccati.cpp:93-111)
    (2, 2, 0, 0, 8, 5, 1e-14, "explicit"),
    (12, 4, 0, 0, 50, 32, 1e-11, "ref"),     # C2 shape
    (36, 12, 0, 0, 100, 6, 1e-11, "ref"),    # C1 shape, bench mu
    (36, 12, 0, 0, 100, 2, 1e-14, "ref"),    # random_long_problem mu (riccati.cpp:113-121)
    (36, 12, 6, 4, 25, 3, 1e-11, "ref"),     # C3 shape (shorter horizon)
    (3, 2, 0, 1, 100, 4, 1e-12, "ref"),      # parametric (riccati.cpp:160-205)
    (8, 3, 2, 0, 15, 4, 1e-8, "explicit"),
    (7, 5, 3, 2, 9, 4, 1e-6, "ref"),
    (17, 1, 0, 0, 12, 3, 1e-9, "ref"),
    (33, 9, 4, 3, 6, 2, 1e-10, "ref"),
    (5, 0, 2, 0, 6, 2, 1e-6, "ref"),         # nu = 0 everywhere (Z = C / mueq, riccati-kernel.hxx:105-108)
    (48, 16, 0, 0, 8, 2, 1e-11, "ref"),
    (12, 4, 0, 0, 30, 4, 1e-3, "explicit_psd"),  # mudyn |Ptilde| large: LDL^T path of the specialised sweep
    (36, 12, 0, 0, 20, 2, 1e-2, "explicit_psd"),
    (12, 4, 0, 0, 30, 3, 1e-7, "explicit_psd"),
    (36, 12, 6, 4, 12, 2, 1e-3, "explicit_psd"),  # C3 shape, LDL^T (non-Neumann) path with theta columns
    (36, 12, 6, 4, 40, 2, 1e-8, "explicit"),
    (76, 32, 0, 0, 12, 3, 1e-11, "ref"),     # C5 shape (Talos walk): streaming sweep, sweep_big.cu
    (76, 32, 0, 0, 6, 2, 1e-8, "explicit"),
    (76, 32, 0, 0, 5, 2, 1e-2, "explicit_psd"),   # C5 shape, LDL^T (non-Neumann) path
    (12, 4, 0, 0, 1, 3, 1e-11, "ref"),       # horizon 1 on every specialised sweep: one stage knot + terminal
    (36, 12, 0, 0, 1, 2, 1e-11, "ref"),
    (36, 12, 6, 4, 1, 2, 1e-9, "ref"),
    (76, 32, 0, 0, 1, 2, 1e-11, "ref"),
    (12, 4, 0, 0, 0, 2, 1e-11, "ref"),       # horizon 0 with a specialised shape: generic kernel (no stage knot)
    (36, 12, 12, 8, 12, 2, 1e-9, "ref"),     # the widest constrained / parametric instantiation (nc = nu: a quadruped's contacts)
    (36, 12, 12, 8, 8, 2, 1e-3, "explicit_psd"),
]

FAST_DIMS = ((12, 4, 0, 0), (36, 12, 0, 0), (76, 32, 0, 0), (36, 12, 6, 4), (36, 12, 12, 8))   # order of padded_shape_for (pad.cu)
# shapes whose knot record exceeds the generic kernel's shared-memory plan: specialised sweep only
NO_GENERIC_DIMS = ((76, 32, 0, 0),)


@pytest.mark.parametrize("case", CASES, ids=lambda c: "nx%d_nu%d_nc%d_nth%d_T%d_B%d_%g_%s" % c)
def test_prox_parity(capi, oracle, case):
    nx, nu, nc, nth, T, B, mu, dist = case
    _prox_parity(capi, oracle, case, capi.STORE_ALL_VALUES)


GENERIC_OK = [c for c in CASES if (c[0], c[1], c[2], c[3]) not in NO_GENERIC_DIMS]


@pytest.mark.parametrize("case", GENERIC_OK, ids=lambda c: "nx%d_nu%d_nc%d_nth%d_T%d_B%d_%g_%s" % c)
def test_prox_parity_generic_kernel(capi, oracle, case):
    """Every shape forced through the generic (any-dimension) kernel: the ones with a specialised (DMMA)
    sweep, and the ones that would otherwise be embedded in a specialised shape by padding."""
    _prox_parity(capi, oracle, case, capi.STORE_ALL_VALUES | capi.FORCE_GENERIC)


def _covering_shape(nx, nu, nc, nth):
    for s in FAST_DIMS:
        if nx <= s[0] and nu <= s[1] and nc <= s[2] and nth <= s[3]:
            return s
    return None


# shapes WITHOUT an instantiated sweep: solved by the covering instantiated sweep through
# exact zero / identity padding (csrc/pad.cu) -- same parity bar, and the specialised sweep must have run
OFF_GRID = [
    # nx, nu, nc, nth, T, batch, mu, dist
    (14, 7, 0, 0, 20, 3, 1e-10, "ref"),          # -> (36, 12, 0, 0)
    (10, 3, 0, 0, 30, 4, 1e-11, "ref"),          # -> (12, 4, 0, 0)
    (12, 4, 2, 0, 15, 3, 1e-9, "ref"),           # constraints -> (36, 12, 6, 4)
    (36, 12, 6, 2, 10, 2, 1e-9, "ref"),          # fewer theta columns -> (36, 12, 6, 4)
    (36, 12, 3, 0, 12, 2, 1e-8, "ref"),
    (20, 8, 3, 1, 12, 3, 1e-9, "ref"),
    (50, 20, 0, 0, 8, 2, 1e-11, "ref"),          # -> (76, 32, 0, 0)
    (40, 12, 0, 0, 10, 2, 1e-11, "ref"),
    (64, 24, 0, 0, 6, 2, 1e-10, "ref"),
    (29, 11, 0, 0, 25, 2, 1e-11, "ref"),         # Solo-12-like (nq 18 + nv 18 minus base?): odd sizes
    (6, 2, 0, 0, 40, 5, 1e-12, "explicit_psd"),  # explicit dynamics survive the embedding (E' = -I)
    (24, 6, 0, 0, 15, 2, 1e-3, "explicit_psd"),  # LDL^T (non-Neumann) path
    (5, 0, 2, 0, 6, 2, 1e-6, "ref"),             # nu = 0
    (3, 2, 0, 1, 100, 4, 1e-12, "ref"),          # parametric, long horizon
    (36, 12, 10, 0, 10, 2, 1e-9, "ref"),         # 6 < nc <= 12 -> (36, 12, 12, 8)
    (36, 12, 6, 8, 10, 2, 1e-9, "ref"),          # more theta columns
    (30, 10, 8, 6, 12, 2, 1e-8, "ref"),
]


@pytest.mark.parametrize("case", OFF_GRID, ids=lambda c: "nx%d_nu%d_nc%d_nth%d_T%d_B%d_%g_%s" % c)
def test_off_grid_shapes_run_the_specialised_sweep(capi, oracle, case):
    nx, nu, nc, nth, T, B, mu, dist = case
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=7 * nx + T, dist=dist)
    s = capi.Solver.for_problems(pp, flags=capi.STORE_ALL_VALUES)
    assert s.padded_shape() == _covering_shape(nx, nu, nc, nth)
    s.upload(pp); s.backward(mu, mu)
    assert s.used_fast_path()
    assert s.used_explicit_path() == dist.startswith("explicit")
    s.close()
    _prox_parity(capi, oracle, case, capi.STORE_ALL_VALUES)      # trajectories, gains, value, kkt0, KKT kernel
    # and without padding the same shape is the generic kernel's: both answers agree to round-off
    try:
        s2 = capi.Solver.for_problems(pp, flags=capi.NO_PADDING)
    except capi.GarB200Error:
        assert nx > 48          # beyond the generic kernel's shared-memory plan: padding is the only path
        return
    s1 = capi.Solver.for_problems(pp)
    assert s2.padded_shape() is None
    outs = []
    theta = np.random.default_rng(1).standard_normal((B, nth)) if nth else None
    for s in (s1, s2):
        s.upload(pp); s.backward(mu, mu); outs.append(s.forward(theta)); s.close()
    for a, b in zip(*outs):
        assert relerr(a, b) < 2 * RTOL       # each is within RTOL of the oracle (checked above)


def test_ragged_knots_on_a_native_shape_are_padded(capi, oracle):
    """Per-knot nu_t / nc_t below the maxima (a native (36, 12, 6, 4) problem whose knots differ): embedded in the
    uniform shape instead of falling back to the generic kernel."""
    rng = np.random.default_rng(3)
    T, B, mu = 14, 2, 1e-9
    nu_t = [int(v) for v in rng.integers(7, 13, T)] + [0]
    nc_t = [int(v) for v in rng.integers(0, 4, T)] + [0]        # (fewer equality rows than controls: well posed)
    nu_t[3], nc_t[3] = 12, 6
    full = generate_problems(B, T, 36, 12, 6, 4, seed=11)
    L = full.layout
    knots = np.zeros_like(full.knots)
    for b in range(B):
        for t in range(T + 1):
            f = L.unpack(full.knots[b, t], int(full.nu_t[t]), int(full.nc_t[t]))
            m, c = nu_t[t], nc_t[t]
            cut = {"S": f["S"][:, :m], "R": f["R"][:m, :m], "r": f["r"][:m], "B": f["B"][:, :m], "C": f["C"][:c],
                   "D": f["D"][:c, :m], "d": f["d"][:c], "Gu": f["Gu"][:m]}
            L.pack(knots[b, t], {**f, **cut}, m, c)
    pp = type(full)(L, T, full.nc0, np.array(nu_t, np.int32), np.array(nc_t, np.int32), knots, full.G0, full.g0)
    theta = rng.standard_normal((B, 4))
    s = capi.Solver.for_problems(pp)
    assert s.padded_shape() == (36, 12, 6, 4)
    s.upload(pp); s.backward(mu, mu)
    assert s.used_fast_path()
    xs, us, vs, lb = s.forward(theta)
    G = GainLayout(36, 12, 6, 4)
    gains = s.download_gains()
    sg = capi.Solver.for_problems(pp, flags=capi.NO_PADDING)      # the generic kernel on the same ragged problem
    sg.upload(pp); sg.backward(mu, mu)
    assert not sg.used_fast_path()
    gen = sg.forward(theta)
    sg.close()
    for b in range(B):
        ref = oracle.prox_solve(pp, mu, mu, theta=theta[b], b=b)
        for got, g, key in ((xs[b], gen[0][b], "xs"), (us[b], gen[1][b], "us"), (vs[b], gen[2][b], "vs"), (lb[b], gen[3][b], "lbdas")):
            # (random ragged equality rows make this problem less well conditioned than the uniform cases: the bar
            #  is the generic kernel's own distance from the oracle on it)
            assert relerr(got, ref[key]) < max(RTOL, 4 * relerr(g, ref[key])), (b, key, relerr(got, ref[key]), relerr(g, ref[key]))
        for t in range(T + 1):
            a = G.unpack(gains[b, t], nu_t[t], nc_t[t]); r = G.unpack(ref["gains"][t], nu_t[t], nc_t[t])
            scale = max(np.abs(r[1]).max() if r[1].size else 0.0, np.abs(r[0]).max() if r[0].size else 0.0, 1e-300)
            for x, y in zip(a, r):
                if x.size:
                    assert np.abs(x - y).max() / scale < RTOL_GAINS, (b, t)
    s.close()


def test_padded_solver_reembeds_only_a_changed_problem(capi, oracle):
    """Re-solving an unchanged problem on a padded shape does not re-embed it; a new upload costs two
    more launches (the cycleAppend path on padded shapes is in test_cycle_append)."""
    nx, nu, T, B, mu = 10, 3, 12, 2, 1e-10
    pp = generate_problems(B, T, nx, nu, seed=5)
    s = capi.Solver.for_problems(pp)
    s.upload(pp); s.backward(mu, mu)
    x1 = s.forward()
    n0 = s.kernel_launches()
    s.backward(mu, mu); x2 = s.forward()
    n1 = s.kernel_launches()
    for a, b in zip(x1, x2):
        assert np.array_equal(a, b)
    s.upload(pp); s.backward(mu, mu); s.forward()
    assert s.kernel_launches() - n1 == (n1 - n0) + 2        # two more launches: the re-embedding and the E == -I check
    s.close()


def _prox_parity(capi, oracle, case, flags):
    nx, nu, nc, nth, T, B, mu, dist = case
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=nx * 131 + T, dist=dist)
    theta = np.random.default_rng(1).standard_normal((B, nth)) if nth else None
    s = capi.Solver.for_problems(pp, flags=flags)
    s.upload(pp)
    s.backward(mu, mu)
    fast_expected = (_covering_shape(nx, nu, nc, nth) is not None and not (flags & (capi.FORCE_GENERIC | capi.NO_PADDING))
                     and T >= 1)
    assert s.used_fast_path() == fast_expected
    xs, us, vs, lb = s.forward(theta)
    assert (s.get_flags() == 0).all()
    G = GainLayout(nx, nu, nc, nth)
    gains = s.download_gains()
    kkt_gpu = s.kkt_error(mu, mu, theta)
    for b in range(B):
        th = None if theta is None else theta[b]
        ref = oracle.prox_solve(pp, mu, mu, theta=th, b=b)
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            assert relerr(got, ref[key]) < RTOL, (b, key, relerr(got, ref[key]))
        # gains, knot by knot (ff / fb / fth with the knot's own row count)
        for t in range(T + 1):
            a = G.unpack(gains[b, t], pp.nu_t[t], pp.nc_t[t])
            r = G.unpack(ref["gains"][t], pp.nu_t[t], pp.nc_t[t])
            scale = max(np.abs(r[1]).max() if r[1].size else 0.0, np.abs(r[0]).max() if r[0].size else 0.0, 1e-300)
            for x, y in zip(a, r):
                if x.size:
                    assert np.abs(x - y).max() / scale < RTOL_GAINS, (b, t)
        # value function + initial stage
        v0 = s.get_value(b, 0)
        from aligator_swk_b200.layout import ValueLayout
        rv0 = ValueLayout(nx, nth).unpack(ref["values"][0])
        assert relerr(v0["Pmat"], rv0["Pmat"]) < RTOL_GAINS and relerr(v0["pvec"], rv0["pvec"]) < RTOL_GAINS
        ff0, fth0, thg, thh = s.get_initial(b)
        assert relerr(ff0, ref["kkt0_ff"]) < RTOL
        if nth:
            assert relerr(fth0, ref["kkt0_fth"]) < RTOL_GAINS
            assert relerr(thg, ref["thGrad"]) < RTOL_GAINS and relerr(thh, ref["thHess"]) < RTOL_GAINS
        # KKT residual kernel == oracle's lqrComputeKktError on the GPU trajectory
        ek = oracle.kkt_error(pp, xs[b], us[b], vs[b], lb[b], mu, mu, theta=th, b=b)
        # (residuals live at round-off level: summation order differs, so compare up to the
        #  rounding scale |data| * |trajectory| * eps * terms, not bit for bit)
        floor = 64 * np.finfo(float).eps * max(1.0, np.abs(pp.knots[b]).max()) * max(
            1.0, np.abs(xs[b]).max(), np.abs(lb[b]).max(), np.abs(us[b]).max() if us[b].size else 0.0)
        assert np.allclose(kkt_gpu[b], ek, rtol=1e-6, atol=floor), (kkt_gpu[b], ek, floor)
        # and it is as small as the oracle's own residual (same order of magnitude)
        assert kkt_gpu[b].max() <= max(10 * ref["kkt"].max(), 1e-12, 10 * floor)
    s.close()


def test_reference_test_tolerances(capi, oracle):
    """random_long_problem (same knot for all t, mu = 1e-14): KKT <= 1e-9 (riccati.cpp:113-135)."""
    pp = generate_problems(2, 100, 36, 12, seed=42, same_knot=True, x0=np.zeros(36))
    s, xs, us, vs, lb = solve_gpu(capi, pp, 1e-14)
    assert s.kkt_error(0.0, 0.0).max() <= 1e-9
    compare_traj(oracle, pp, 1e-14, xs, us, vs, lb)


def test_ragged_knots_short_horz_pb(capi, oracle):
    """riccati.cpp:20-91: one knot with nc = nu among nc = 0 knots, E = -I."""
    from test_oracle import short_horz_pb
    pp = short_horz_pb().pack()
    s, xs, us, vs, lb = solve_gpu(capi, pp, 1e-14)
    assert s.kkt_error(0.0, 0.0).max() <= 1e-9
    compare_traj(oracle, pp, 1e-14, xs, us, vs, lb)


def test_golden_fixtures(capi):
    gd = os.path.join(os.path.dirname(__file__), "golden")
    from golden.make_golden import load_case
    for f in sorted(x for x in os.listdir(gd) if x.endswith(".npz")):
        pp, mu, theta, gold = load_case(os.path.join(gd, f))
        s, xs, us, vs, lb = solve_gpu(capi, pp, mu, theta)
        for got, key in ((xs[0], "xs"), (us[0], "us"), (vs[0], "vs"), (lb[0], "lbdas")):
            assert relerr(got, gold[key]) < RTOL, (f, key, relerr(got, gold[key]))


def test_gar_mirror_api(capi, oracle):
    """The reference-shaped host API (gar.ProximalRiccatiSolver) end to end, incl. getFeedback."""
    from aligator_swk_b200 import gar
    from test_oracle import short_horz_pb
    prob = short_horz_pb()
    solver = gar.ProximalRiccatiSolver(prob)
    assert solver.backward(1e-14, 1e-14)
    xs, us, vs, lbdas = gar.lqrInitializeSolution(prob)
    assert len(xs) == prob.horizon + 1 and len(us) == prob.horizon + 1 and len(lbdas) == prob.horizon + 1
    assert solver.forward(xs, us, vs, lbdas)
    err = gar.lqrComputeKktError(prob, xs, us, vs, lbdas, 0.0, 0.0)
    assert err.max() <= 1e-9
    K = solver.getFeedback(0)
    kff = solver.getFeedforward(0)
    assert K.shape == (2 + 0 + 4, 2) and kff.shape == (6,)
    np.testing.assert_allclose(us[0], kff[:2] + K[:2] @ xs[0], rtol=1e-12, atol=1e-14)
    # the solver re-reads the problem on every backward (proximal-riccati.hxx:37)
    prob.stages[1].q[:] = 5.0
    solver.backward(1e-14, 1e-14)
    xs2, us2, vs2, lb2 = gar.lqrInitializeSolution(prob)
    solver.forward(xs2, us2, vs2, lb2)
    assert np.abs(us2[0] - us[0]).max() > 1e-6
    ref = oracle.prox_solve(prob.pack(), 1e-14, 1e-14)
    assert relerr(np.array(xs2), ref["xs"]) < RTOL


def test_gar_mirror_dense_solver(capi, oracle):
    """RiccatiSolverDense (dense-riccati.hpp:28-58) through the host mirror: same API and the same solution as
    the proximal solver (tests/gar/riccati.cpp:143-157) — from a different kernel (csrc/dense.cu), compared
    with the oracle's own restatement of the dense kernel; a non-zero Gv is READ (dense-kernel.hpp:76,134)."""
    from aligator_swk_b200 import gar
    from test_oracle import short_horz_pb
    prob = short_horz_pb()
    dense = gar.RiccatiSolverDense(prob)
    assert dense.backward(1e-14, 1e-14)
    assert not dense._solver.used_fast_path()
    xs, us, vs, lbdas = gar.lqrInitializeSolution(prob)
    assert dense.forward(xs, us, vs, lbdas)
    assert gar.lqrComputeKktError(prob, xs, us, vs, lbdas, 0.0, 0.0).max() <= 1e-9
    ref = oracle.dense_solve(prob.pack(), 1e-14, 1e-14)
    assert relerr(np.array(xs), ref["xs"]) < RTOL
    assert relerr(np.array(xs), oracle.prox_solve(prob.pack(), 1e-14, 1e-14)["xs"]) < 1e-9
    # Gv != 0 changes the answer of the dense solver and only of the dense solver
    prob.addParameterization(1)
    rng = np.random.default_rng(3)
    for k in prob.stages:
        k.Gx[:] = rng.standard_normal(k.Gx.shape); k.Gu[:] = rng.standard_normal(k.Gu.shape)
        k.Gth[:] = 1.0; k.gamma[:] = rng.standard_normal(1)
        if k.nc:
            k.Gv[:] = rng.standard_normal(k.Gv.shape)
    theta = np.array([0.7])
    dense = gar.RiccatiSolverDense(prob)
    assert dense.backward(1e-10, 1e-10)
    xs, us, vs, lbdas = gar.lqrInitializeSolution(prob)
    dense.forward(xs, us, vs, lbdas, theta)
    pp = prob.pack()
    gv = np.zeros((len(prob.stages), 1, pp.layout.nc))
    for t, k in enumerate(prob.stages):
        gv[t, :, :k.nc] = np.asarray(k.Gv).reshape(k.nc, 1).T
    ref = oracle.dense_solve(pp, 1e-10, 1e-10, theta=theta, gv=gv)
    ref0 = oracle.dense_solve(pp, 1e-10, 1e-10, theta=theta, gv=None)
    assert relerr(ref["vs"], ref0["vs"]) > 1e-6                      # Gv matters
    assert relerr(np.array(xs), ref["xs"]) < RTOL
    U = np.zeros_like(ref["us"])
    for t, u in enumerate(us):
        U[t, :len(u)] = u
    assert relerr(U, ref["us"]) < RTOL
    assert relerr(dense.thGrad, ref["thGrad"]) < 1e-9 and relerr(dense.thHess, ref["thHess"]) < 1e-9


DENSE_CASES = [
    # nx, nu, nc, nth, T, B, mu, dist
    (12, 4, 0, 0, 50, 3, 1e-11, "ref"),
    (36, 12, 0, 0, 100, 2, 1e-11, "ref"),          # C1 shape at its own size
    (36, 12, 6, 4, 30, 2, 1e-9, "ref"),
    (7, 5, 3, 2, 9, 3, 1e-6, "ref"),
    (5, 0, 2, 0, 6, 2, 1e-6, "ref"),               # nu = 0
    (2, 2, 0, 0, 0, 2, 1e-10, "ref"),              # horizon 0
    (36, 12, 0, 0, 20, 2, 1e-9, "explicit"),       # indefinite stages: what the dense solver is for
]


@pytest.mark.parametrize("dims", [(50, 20, 4, 0, 6, 2), (52, 12, 3, 1, 5, 2)],
                         ids=lambda d: "nx%d_nu%d_nc%d_nth%d_T%d_B%d" % d)
def test_shapes_nothing_else_holds_take_the_dense_kernel(capi, oracle, dims):
    """nx > 36 with constraints / parameters (no covering instantiation, beyond the generic kernel's shared-memory
    plan): garb200_create falls back to the dense per-knot KKT solver on its own -- slow, but nothing that fits it is
    rejected.  Same interface and records; checked against the proximal oracle."""
    nx, nu, nc, nth, T, B = dims
    mu = 1e-8
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=3 * nx + T)
    theta = np.random.default_rng(4).standard_normal((B, nth)) if nth else None
    s = capi.Solver.for_problems(pp)
    assert s.padded_shape() is None
    s.upload(pp); s.backward(mu, mu)
    assert not s.used_fast_path()
    xs, us, vs, lb = s.forward(theta)
    assert (s.get_flags() == 0).all()
    for b in range(B):
        ref = oracle.prox_solve(pp, mu, mu, theta=None if theta is None else theta[b], b=b)
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            assert relerr(got, ref[key]) < 1e-9, (b, key, relerr(got, ref[key]))
    assert s.kkt_error(mu, mu, theta).max() < 1e-7
    s.close()
    with pytest.raises(capi.GarB200Error):          # ... and what does not fit it still says so
        capi.Solver(76, 32, 6, 0, horizon=4)


@pytest.mark.parametrize("case", DENSE_CASES, ids=lambda c: "nx%d_nu%d_nc%d_nth%d_T%d_B%d_%g_%s" % c)
def test_dense_kernel_parity(capi, oracle, case):
    """csrc/dense.cu against the oracle's restatement of DenseKernel / RiccatiSolverDense (oracle/dense.cpp):
    trajectories, gains knot by knot, value function, kkt0, thGrad / thHess; random non-zero Gv."""
    nx, nu, nc, nth, T, B, mu, dist = case
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=nx + 7 * T, dist=dist)
    rng = np.random.default_rng(2)
    theta = rng.standard_normal((B, nth)) if nth else None
    gv = rng.standard_normal((B, T + 1, nth, nc)) if nc * nth else None
    s = capi.Solver.for_problems(pp, flags=capi.DENSE_KERNEL | capi.STORE_ALL_VALUES)
    s.upload(pp)
    if gv is not None:
        s.upload_gv(gv)
    s.backward(mu, mu)
    assert not s.used_fast_path()
    xs, us, vs, lb = s.forward(theta)
    assert (s.get_flags() == 0).all()
    G = GainLayout(nx, nu, nc, nth)
    gains = s.download_gains()
    from aligator_swk_b200.layout import ValueLayout
    for b in range(B):
        th = None if theta is None else theta[b]
        ref = oracle.dense_solve(pp, mu, mu, theta=th, gv=None if gv is None else gv[b], b=b)
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            assert relerr(got, ref[key]) < RTOL, (b, key, relerr(got, ref[key]))
        for t in range(T + 1):
            a = G.unpack(gains[b, t], pp.nu_t[t], pp.nc_t[t])
            r = G.unpack(ref["gains"][t], pp.nu_t[t], pp.nc_t[t])
            scale = max(np.abs(r[1]).max() if r[1].size else 0.0, np.abs(r[0]).max() if r[0].size else 0.0, 1e-300)
            for x, y in zip(a, r):
                if x.size:
                    assert np.abs(x - y).max() / scale < RTOL_GAINS, (b, t)
        v0, rv0 = s.get_value(b, 0), ValueLayout(nx, nth).unpack(ref["values"][0])
        assert relerr(v0["Pmat"], rv0["Pmat"]) < RTOL_GAINS and relerr(v0["pvec"], rv0["pvec"]) < RTOL_GAINS
        ff0, fth0, thg, thh = s.get_initial(b)
        assert relerr(ff0, ref["kkt0_ff"]) < RTOL
        if nth:
            assert relerr(fth0, ref["kkt0_fth"]) < RTOL_GAINS
            assert relerr(thg, ref["thGrad"]) < RTOL_GAINS and relerr(thh, ref["thHess"]) < RTOL_GAINS
    s.close()
    with pytest.raises(capi.GarB200Error):           # no parallel-in-time mode, stage matrix must fit shared memory
        capi.Solver(12, 4, horizon=8, num_legs=2, flags=capi.DENSE_KERNEL)
    with pytest.raises(capi.GarB200Error):
        capi.Solver(76, 32, horizon=4, flags=capi.DENSE_KERNEL)


def test_error_conventions(capi):
    with pytest.raises(capi.GarB200Error):
        capi.Solver(4, 2, horizon=5, nc0=6)          # nc0 > nx
    with pytest.raises(capi.GarB200Error):
        capi.Solver(4, 2, nth=1, horizon=5, num_legs=2)  # legs need nth == 0
    s = capi.Solver(4, 2, horizon=5)
    with pytest.raises(capi.GarB200Error) as e:
        s.backward(1e-8, 1e-8)                         # no problem uploaded
    assert e.value.code == capi.ERR_STATE
    pp = generate_problems(1, 5, 4, 2, seed=0)
    s.upload(pp)
    with pytest.raises(capi.GarB200Error):
        s.forward()                                    # forward before backward
    # singular E is flagged, not silently wrong
    pp.knots[0, 2, pp.layout.offsets["E"]: pp.layout.offsets["E"] + 16] = 0.0
    s.upload(pp)
    s.backward(1e-8, 1e-8)
    with pytest.raises(capi.GarB200Error) as e:
        s.synchronize()
    assert e.value.code == capi.ERR_NUMERIC
    assert s.get_flags()[0] & 1
    # nx = 76 only has the streaming sweep: the generic kernel's shared-memory plan does not fit
    with pytest.raises(capi.GarB200Error):
        capi.Solver(76, 32, horizon=5, flags=capi.FORCE_GENERIC)


def test_device_generator_roundtrip(capi, oracle):
    """bench data source: generate on device, download a sample, oracle on those very bytes."""
    from aligator_swk_b200.synth import PackedProblems
    from aligator_swk_b200.layout import KnotLayout
    nx, nu, T, B, mu = 12, 4, 50, 300, 1e-11
    s = capi.Solver(nx, nu, horizon=T, batch=B)
    s.generate(seed=42, dist=0)
    s.backward(mu, mu)
    xs, us, vs, lb = s.forward()
    assert (s.get_flags() == 0).all()
    kkt = s.kkt_error(mu, mu)
    assert kkt.max() < 1e-4      # E ~ 1000 N(0,1), f ~ N(0, 100^2): residual scale of the data itself
    idx = [0, 1, 149, 299, int(kkt.max(axis=1).argmax())]
    for b in idx:
        knots, G0, g0 = s.download_packed(b, 1)
        pp = PackedProblems(KnotLayout(nx, nu, 0, 0), T, nx, s.nu_t, s.nc_t, knots, G0, g0)
        ref = oracle.prox_solve(pp, mu, mu)
        assert relerr(xs[b], ref["xs"]) < RTOL and relerr(us[b], ref["us"]) < RTOL
        assert relerr(lb[b], ref["lbdas"]) < RTOL
        assert kkt[b].max() <= max(10 * ref["kkt"].max(), 1e-10)   # as small as the CPU solver's own
    # distribution sanity (test_util.cpp:31-35): E ~ 1000 N(0,1), B ~ U[-1,1]
    knots, _, _ = s.download_packed(0, 32)
    o = KnotLayout(nx, nu, 0, 0).offsets
    E = knots[:, :T, o["E"]: o["E"] + nx * nx]
    Bm = knots[:, :T, o["B"]: o["B"] + nx * nu]
    assert 900 < E.std() < 1100 and abs(E.mean()) < 20
    assert Bm.min() >= -1 and Bm.max() <= 1 and 0.55 < Bm.std() < 0.6


LEG_CASES = [
    # nx, nu, nc, T, legs, batch, mu
    (32, 12, 0, 50, 6, 2, 1e-9),      # parallel_solver_class (tests/gar/parallel.cpp:199-271)
    (4, 2, 0, 11, 2, 3, 1e-9),
    (12, 4, 0, 50, 5, 4, 1e-10),
    (36, 12, 0, 127, 8, 1, 1e-9),     # specialised leg sweep + specialised condensed solve
    (36, 12, 0, 50, 6, 3, 1e-9),
    (36, 12, 0, 21, 2, 2, 1e-11),
    (36, 12, 0, 9, 9, 2, 1e-8),       # one- and two-knot legs on the specialised path
    (8, 3, 2, 20, 3, 2, 1e-8),
    (6, 2, 0, 7, 8, 2, 1e-9),         # one-knot legs
]


@pytest.mark.parametrize("case", LEG_CASES, ids=lambda c: "nx%d_nu%d_nc%d_T%d_J%d_B%d_%g" % c)
def test_parallel_in_time_parity(capi, oracle, case):
    """ParallelRiccatiSolver on the GPU (legs = CTAs): serial == parallel <= 1e-10
    (tests/gar/parallel.cpp:243-260) and == the oracle's restated parallel solver."""
    nx, nu, nc, T, J, B, mu = case
    pp = generate_problems(B, T, nx, nu, nc, 0, seed=nx + T + J, dist="ref")
    s, xs, us, vs, lb = solve_gpu(capi, pp, mu, num_legs=J)
    assert (s.get_flags() == 0).all()
    kkt = s.kkt_error(mu, mu)
    for b in range(B):
        ser = oracle.prox_solve(pp, mu, mu, b=b)
        par = oracle.parallel_solve(pp, mu, mu, J, b=b)
        assert kkt[b].max() <= max(1e-10, 100 * par["kkt"].max()), (kkt[b], par["kkt"])
        scale_x = max(np.abs(ser["xs"]).max(), 1.0)
        scale_l = max(np.abs(ser["lbdas"]).max(), 1.0)
        assert np.abs(xs[b] - ser["xs"]).max() <= 1e-10 * scale_x
        assert np.abs(lb[b] - ser["lbdas"]).max() <= 1e-9 * scale_l
        assert relerr(xs[b], par["xs"]) < 1e-9 and relerr(us[b], par["us"]) < 1e-9
    s.close()


@pytest.mark.parametrize("dims", [(4, 2, 5, 3), (36, 12, 5, 3)], ids=lambda d: "nx%d_nu%d_T%d_J%d" % d)
def test_collapse_feedback(capi, oracle, dims):
    """collapseFeedback (parallel-solver.hpp:55-66): K0 -= Kth0 * subdiagonal[1].  Explicit
    dynamics and two-knot legs so that the correction is not vanishingly small."""
    (nx, nu, T, J), mu = dims, 1e-8
    pp = generate_problems(2, T, nx, nu, 0, 0, seed=11, dist="explicit")
    s, xs, us, vs, lb = solve_gpu(capi, pp, mu, num_legs=J)
    G = GainLayout(nx, nu, 0, nx)
    for b in range(2):
        g_before = s.get_gains(b, 0)[1].copy()
        ref0 = oracle.parallel_solve(pp, mu, mu, J, collapse=False, b=b)
        rfb0 = G.unpack(ref0["gains"][0], nu, 0)[1]
        assert np.abs(g_before - rfb0).max() <= 1e-9 * max(np.abs(rfb0).max(), 1.0)
    s.collapse_feedback()
    for b in range(2):
        g_after = s.get_gains(b, 0)[1]
        ref = oracle.parallel_solve(pp, mu, mu, J, collapse=True, b=b)
        rfb = G.unpack(ref["gains"][0], nu, 0)[1]
        rfb0 = G.unpack(oracle.parallel_solve(pp, mu, mu, J, b=b)["gains"][0], nu, 0)[1]
        assert np.abs(rfb[:nu] - rfb0[:nu]).max() > 1e-6          # the correction is visible
        assert np.abs(g_after - rfb).max() <= 1e-9 * max(np.abs(rfb).max(), 1.0)
    s.close()


@pytest.mark.parametrize("dims", [(4, 2, 6, 3), (12, 4, 7, 3), (36, 12, 5, 2), (76, 32, 4, 2), (50, 20, 5, 2)], ids=lambda d: "nx%d_nu%d_T%d_B%d" % d)
def test_cycle_append(capi, oracle, dims):
    """cycleAppend (proximal-riccati.hxx:80-88, mpc-util.hpp:17-22): factor storage rotates
    left over [0, N), the new knot becomes knot N-1, and the next solve sees the rotated problem.
    The nx = 12 / 36 shapes run the specialised sweep and rollout kernels through the ring offset."""
    nx, nu, T, B = dims
    mu = 1e-10
    pp = generate_problems(B, T, nx, nu, seed=5)
    extra = generate_problems(1, 1, nx, nu, seed=6)
    s = capi.Solver.for_problems(pp)
    s.upload(pp)
    s.backward(mu, mu)
    g0 = s.download_gains()
    L = pp.layout
    rec = np.ascontiguousarray(extra.knots[0, 0])
    o = L.offsets
    view = capi.KnotView()
    view.nu, view.nc = nu, 0
    keep = {}
    for name in ("Q", "S", "R", "q", "r", "A", "B", "E", "f"):
        r, c = L.shapes()[name]
        keep[name] = np.ascontiguousarray(rec[o[name]: o[name] + r * c])
        setattr(view, name, keep[name].ctypes.data_as(capi._dp))
    for rep in range(2):     # twice: the ring offset accumulates
        s.cycle_append([view], shared=True)
        g1 = s.download_gains()
        assert np.array_equal(g1[:, : T - 1], g0[:, 1:T]) and not g1[:, T - 1].any()
        assert np.array_equal(g1[:, T], g0[:, T])
        knots = pp.knots.copy()
        knots[:, : T - 1] = pp.knots[:, 1:T]
        knots[:, T - 1] = rec
        pp = type(pp)(L, T, pp.nc0, pp.nu_t, pp.nc_t, knots, pp.G0, pp.g0)
        k2, _, _ = s.download_packed(0, B)
        assert np.array_equal(k2, knots)
        s.backward(mu, mu)
        xs, us, vs, lb = s.forward()
        compare_traj(oracle, pp, mu, xs, us, vs, lb)
        g0 = s.download_gains()
    # a full re-upload after cycling lands in the rotated slots too
    s.upload(pp)
    s.backward(mu, mu)
    xs, us, vs, lb = s.forward()
    compare_traj(oracle, pp, mu, xs, us, vs, lb)
    s.close()


@pytest.mark.parametrize("shape", [(12, 4, 20), (36, 12, 12), (76, 32, 6)])
def test_fast_path_indefinite_kkt_falls_back_to_bunch_kaufman(capi, oracle, shape):
    """The specialised sweep factors the reduced KKT matrix without pivoting after checking its
    inertia; a knot with an indefinite Rhat must be re-factorised with Bunch-Kaufman pivoting and
    still match the oracle.  Problems 0 and 2 are untouched, problem 1 has R_t = -R_t at t = 3."""
    nx, nu, T = shape
    mu = 1e-10
    pp = generate_problems(3, T, nx, nu, seed=77)
    o = pp.layout.offsets
    pp.knots[1, 3, o["R"]: o["R"] + nu * nu] *= -1.0
    s, xs, us, vs, lb = solve_gpu(capi, pp, mu)
    assert s.used_fast_path()
    assert (s.get_flags() == 0).all()
    compare_traj(oracle, pp, mu, xs, us, vs, lb, rtol=1e-9)
    s.close()


@pytest.mark.parametrize("shape", [(12, 4, 15), (36, 12, 10), (76, 32, 4)])
def test_fast_path_general_initial_constraint(capi, oracle, shape):
    """The specialised sweep solves the initial stage in reduced form when G0 == I exactly and
    with Bunch-Kaufman on [P0 G0^T; G0 -mu I] otherwise (proximal-riccati.hxx:44-62): a dense
    random G0 must take the second route and still match the oracle."""
    nx, nu, T = shape
    mu = 1e-9
    pp = generate_problems(3, T, nx, nu, seed=5)
    rng = np.random.default_rng(9)
    pp.G0[1] = (np.eye(nx) + 0.3 * rng.standard_normal((nx, nx))).reshape(-1, order="F")
    pp.G0[2] = rng.standard_normal((nx, nx)).reshape(-1, order="F")
    s, xs, us, vs, lb = solve_gpu(capi, pp, mu)
    assert s.used_fast_path() and (s.get_flags() == 0).all()
    compare_traj(oracle, pp, mu, xs, us, vs, lb, rtol=1e-9)
    for b in range(3):
        ref = oracle.prox_solve(pp, mu, mu, b=b)
        assert relerr(s.get_initial(b)[0], ref["kkt0_ff"]) < 1e-9
    s.close()


def test_fast_path_general_initial_constraint_with_theta(capi, oracle):
    """Same for the parametric C3 shape: dense G0 -> CTA Bunch-Kaufman with the theta columns,
    kkt0.fth / thGrad / thHess against the oracle (proximal-riccati.hxx:53-61)."""
    nx, nu, nc, nth, T, mu = 36, 12, 6, 4, 6, 1e-9
    pp = generate_problems(2, T, nx, nu, nc, nth, seed=11, dist="explicit")
    rng = np.random.default_rng(4)
    pp.G0[1] = (np.eye(nx) + 0.3 * rng.standard_normal((nx, nx))).reshape(-1, order="F")
    theta = rng.standard_normal((2, nth))
    s, xs, us, vs, lb = solve_gpu(capi, pp, mu, theta=theta)
    assert s.used_fast_path() and (s.get_flags() == 0).all()
    for b in range(2):
        ref = oracle.prox_solve(pp, mu, mu, theta=theta[b], b=b)
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            assert relerr(got, ref[key]) < 1e-9, (b, key)
        ff0, fth0, thg, thh = s.get_initial(b)
        assert relerr(ff0, ref["kkt0_ff"]) < 1e-9 and relerr(fth0, ref["kkt0_fth"]) < 1e-9
        assert relerr(thg, ref["thGrad"]) < 1e-9 and relerr(thh, ref["thHess"]) < 1e-9
    s.close()
