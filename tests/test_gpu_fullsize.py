"""GPU parity at the BASELINE configs' OWN sizes: every config's shape and
horizon against the CPU oracle on the same bytes, small batches (the batch axis is embarrassingly
parallel; the horizon is what accumulates rounding).  Reference pattern for C4:
/root/reference/tests/gar/parallel.cpp:243-260 (serial == parallel, inf-norm <= 1e-10 * scale).
"""
import numpy as np
import pytest

from aligator_swk_b200.synth import generate_problems
from conftest import relerr

pytestmark = pytest.mark.gpu

RTOL = 1e-10


@pytest.fixture(scope="module")
def capi():
    from aligator_swk_b200 import capi as c
    c.lib()
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return c


def _solve(capi, pp, mu, theta=None, num_legs=0, **tune):
    s = capi.Solver.for_problems(pp, num_legs=num_legs)
    if tune:
        s.set_tuning(**tune)
    s.upload(pp)
    s.backward(mu, mu)
    xs, us, vs, lb = s.forward(theta)
    assert (s.get_flags() == 0).all()
    return s, xs, us, vs, lb


def _check(oracle, pp, mu, xs, us, vs, lb, theta=None, rtol=RTOL):
    worst = 0.0
    for b in range(pp.batch):
        th = None if theta is None else theta[b]
        ref = oracle.prox_solve(pp, mu, mu, theta=th, b=b)
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            e = relerr(got, ref[key])
            worst = max(worst, e)
            assert e < rtol, (b, key, e)
    return worst


@pytest.mark.parametrize("batch", [1, 2, 7], ids=lambda b: "B%d" % b)
def test_c1_full_size_lockstep_groups(capi, oracle, batch):
    """C1 / c1b: nx=36 nu=12 T=100, mu = 1e-11, default launch (one CTA per problem) vs the oracle; the
    opt-in lockstep launch (3 problems per CTA; batches that are not a multiple of 3 pad with a group that
    re-solves the last problem) must give the same bits."""
    pp = generate_problems(batch, 100, 36, 12, seed=100 + batch)
    s, xs, us, vs, lb = _solve(capi, pp, 1e-11)
    assert s.used_fast_path()
    _check(oracle, pp, 1e-11, xs, us, vs, lb)
    g3 = s.download_gains()
    s1, xs1, us1, vs1, lb1 = _solve(capi, pp, 1e-11, group=3)
    assert np.array_equal(xs, xs1) and np.array_equal(us, us1) and np.array_equal(lb, lb1)
    assert np.array_equal(g3, s1.download_gains())
    s.close(); s1.close()


def test_c3_full_size(capi, oracle):
    """C3: nx=36 nu=12 nc=6 nth=4, T=200 (the config's horizon), batch 4, fixed theta ~ N(0,1)."""
    nx, nu, nc, nth, T, B, mu = 36, 12, 6, 4, 200, 4, 1e-11
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=303)
    theta = np.random.default_rng(3).standard_normal((B, nth))
    s, xs, us, vs, lb = _solve(capi, pp, mu, theta)
    assert s.used_fast_path()
    _check(oracle, pp, mu, xs, us, vs, lb, theta)
    kkt = s.kkt_error(mu, mu, theta)
    for b in range(B):
        ref = oracle.prox_solve(pp, mu, mu, theta=theta[b], b=b)
        assert kkt[b].max() <= max(10 * ref["kkt"].max(), 1e-10)
    s.close()


def test_c5_full_size(capi, oracle):
    """C5: nx=76 nu=32, T=100 (the config's horizon), batch 2 — streaming sweep."""
    nx, nu, T, B, mu = 76, 32, 100, 2, 1e-11
    pp = generate_problems(B, T, nx, nu, seed=505)
    s, xs, us, vs, lb = _solve(capi, pp, mu)
    assert s.used_fast_path()
    _check(oracle, pp, mu, xs, us, vs, lb)
    kkt = s.kkt_error(mu, mu)
    for b in range(B):
        ref = oracle.prox_solve(pp, mu, mu, b=b)
        assert kkt[b].max() <= max(10 * ref["kkt"].max(), 1e-10)
    s.close()


def test_c4_full_size_parallel_in_time(capi, oracle):
    """C4: nx=36 nu=12, T=4096 split into 64 legs (128-block condensed chain), one problem, against the
    SERIAL oracle: |dx|_inf, |dlambda|_inf <= 1e-10 * scale (tests/gar/parallel.cpp:243-260)."""
    nx, nu, T, J, mu = 36, 12, 4096, 64, 1e-11
    pp = generate_problems(1, T, nx, nu, seed=4096)
    s, xs, us, vs, lb = _solve(capi, pp, mu, num_legs=J)
    assert s.used_fast_path()
    ser = oracle.prox_solve(pp, mu, mu, b=0)
    sx, sl = max(np.abs(ser["xs"]).max(), 1.0), max(np.abs(ser["lbdas"]).max(), 1.0)
    ex, el = np.abs(xs[0] - ser["xs"]).max() / sx, np.abs(lb[0] - ser["lbdas"]).max() / sl
    eu = np.abs(us[0] - ser["us"]).max() / max(np.abs(ser["us"]).max(), 1.0)
    kkt = s.kkt_error(mu, mu)[0]
    print(f"C4 full size: dx {ex:.2e} dlambda {el:.2e} du {eu:.2e} kkt gpu {kkt} serial-oracle kkt {ser['kkt']} "
          f"refinement steps {s.get_refinement_steps()}")
    assert ex <= 1e-10 and el <= 1e-10 and eu <= 1e-10
    assert kkt.max() <= max(100 * ser["kkt"].max(), 1e-9)
    s.close()


def test_c2_shape_full_horizon_batch(capi, oracle):
    """C2: nx=12 nu=4 T=50 at a batch that spans several CTA waves; every 97th problem vs the oracle."""
    nx, nu, T, B, mu = 12, 4, 50, 4096, 1e-11
    pp = generate_problems(B, T, nx, nu, seed=212)
    s, xs, us, vs, lb = _solve(capi, pp, mu)
    for b in range(0, B, 97):
        ref = oracle.prox_solve(pp, mu, mu, b=b)
        assert relerr(xs[b], ref["xs"]) < RTOL and relerr(us[b], ref["us"]) < RTOL and relerr(lb[b], ref["lbdas"]) < RTOL
    s.close()


def test_cycle_append_then_stage_initial_commit(capi, oracle):
    """ADVICE r1 (capi.cu commit): the natural MPC flow cycle_append(new knot) -> stage_initial(new g0)
    -> commit must not overwrite the appended knot (nor any knot installed by upload_packed) with stale
    staging data: commit uploads only what was staged since the last commit."""
    nx, nu, T, B, mu = 12, 4, 8, 3, 1e-10
    pp = generate_problems(B, T, nx, nu, seed=21)
    extra = generate_problems(1, 1, nx, nu, seed=22)
    L = pp.layout
    o = L.offsets
    s = capi.Solver.for_problems(pp)
    s.upload(pp)                       # installed by upload_packed: the staging mirror never saw it
    rec = np.ascontiguousarray(extra.knots[0, 0])
    view = capi.KnotView()
    view.nu, view.nc = nu, 0
    keep = {}
    for name in ("Q", "S", "R", "q", "r", "A", "B", "E", "f"):
        r, c = L.shapes()[name]
        keep[name] = np.ascontiguousarray(rec[o[name]: o[name] + r * c])
        setattr(view, name, keep[name].ctypes.data_as(capi._dp))
    s.cycle_append([view], shared=True)
    new_g0 = np.random.default_rng(5).standard_normal((B, nx))
    for b in range(B):
        s.stage_initial(b, np.ascontiguousarray(pp.G0[b]), np.ascontiguousarray(new_g0[b]))
    s.commit()
    knots = pp.knots.copy()
    knots[:, : T - 1] = pp.knots[:, 1:T]
    knots[:, T - 1] = rec
    k2, G02, g02 = s.download_packed(0, B)
    assert np.array_equal(k2, knots), "commit overwrote knots it did not stage"
    assert np.array_equal(g02, new_g0)
    pp2 = type(pp)(L, T, pp.nc0, pp.nu_t, pp.nc_t, knots, pp.G0, new_g0)
    s.backward(mu, mu)
    xs, us, vs, lb = s.forward()
    _check(oracle, pp2, mu, xs, us, vs, lb)
    # staging ONE knot of ONE problem and committing touches nothing else either
    view2 = capi.KnotView()
    view2.nu, view2.nc = nu, 0
    rec2 = np.ascontiguousarray(generate_problems(1, 1, nx, nu, seed=23).knots[0, 0])
    keep2 = {}
    for name in ("Q", "S", "R", "q", "r", "A", "B", "E", "f"):
        r, c = L.shapes()[name]
        keep2[name] = np.ascontiguousarray(rec2[o[name]: o[name] + r * c])
        setattr(view2, name, keep2[name].ctypes.data_as(capi._dp))
    s.stage_knot(1, 2, view2)
    s.commit()
    knots[1, 2] = rec2
    k3, _, _ = s.download_packed(0, B)
    assert np.array_equal(k3, knots)
    s.close()


def test_forward_is_const_repeated_theta(capi, oracle):
    """forward() is const in the reference (riccati-base.hpp:20-23): one backward, several forwards with
    different theta, each matching the oracle."""
    nx, nu, nc, nth, T, B, mu = 36, 12, 6, 4, 12, 2, 1e-9
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=77, dist="explicit")
    s = capi.Solver.for_problems(pp)
    s.upload(pp)
    s.backward(mu, mu)
    rng = np.random.default_rng(0)
    for rep in range(3):
        theta = rng.standard_normal((B, nth)) * (rep + 1)
        xs, us, vs, lb = s.forward(theta)
        _check(oracle, pp, mu, xs, us, vs, lb, theta, rtol=1e-9)
    s.close()


def test_tiny_positive_pivot_falls_back_to_bunch_kaufman(capi, oracle):
    """ADVICE r1 / VERDICT 2(e): a rank-deficient PSD Rhat whose control is determined only through
    the constraint rows D.  R(0,0) = 1e-18 with S(:,0) = B(:,0) = 0 gives a tiny POSITIVE first
    pivot: it passes the inertia check, and the unpivoted LDL^T would divide by rounding noise.  The
    rank guard of ldlt_factor_warp must hand the knot to Bunch-Kaufman (which pivots with the D rows,
    bunchkaufman.hpp:20-167) and the result must match the oracle."""
    nx, nu, nc, nth, T, mu = 36, 12, 6, 4, 10, 1e-9
    pp = generate_problems(3, T, nx, nu, nc, nth, seed=91, dist="explicit")
    o = pp.layout.offsets
    k = pp.knots[1, 4]
    R = k[o["R"]: o["R"] + nu * nu].reshape(nu, nu, order="F")
    R[0, :] = 0.0; R[:, 0] = 0.0; R[0, 0] = 1e-18
    k[o["R"]: o["R"] + nu * nu] = R.reshape(-1, order="F")
    k[o["S"]: o["S"] + nx] = 0.0                       # S(:, 0)
    k[o["B"]: o["B"] + nx] = 0.0                       # B(:, 0)
    k[o["Gu"]: o["Gu"] + nu * nth].reshape(nu, nth, order="F")[0, :] = 0.0
    theta = np.random.default_rng(2).standard_normal((3, nth))
    s, xs, us, vs, lb = _solve(capi, pp, mu, theta)
    assert s.used_fast_path()
    _check(oracle, pp, mu, xs, us, vs, lb, theta, rtol=1e-9)
    s.close()


@pytest.mark.parametrize("shape", [(12, 4, 0, 0, 50, 2), (36, 12, 0, 0, 100, 2), (36, 12, 6, 4, 30, 2), (76, 32, 0, 0, 25, 2)],
                         ids=lambda c: "nx%d_nu%d_nc%d_nth%d_T%d_B%d" % c)
def test_explicit_dynamics_path_is_exact(capi, oracle, shape):
    """E_t == -I at every knot (explicit-dynamics.hxx:32, solver-proxddp.hxx:826 — what ProxDDP feeds): detected
    on the device, the sweep then skips the E^-1 pre-pass and never reads E.  Same bits as the general path
    on the same input, and parity with the oracle (batch 2 of this seed is well conditioned: on the long
    explicit-dynamics horizons some problems put the oracle itself 1e-9 from the extended-precision
    solution, see tests/test_reference_fixtures.py); one knot with E != -I switches it off again; the
    GARB200_EXPLICIT_DYNAMICS promise flag takes the same path without the check."""
    nx, nu, nc, nth, T, B = shape
    mu = 1e-9
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=61, dist="explicit_psd")
    theta = np.random.default_rng(6).standard_normal((B, nth)) if nth else None
    s, xs, us, vs, lb = _solve(capi, pp, mu, theta)
    assert s.used_fast_path() and s.used_explicit_path()
    _check(oracle, pp, mu, xs, us, vs, lb, theta)
    g = s.download_gains()
    # general path on the same bytes: identical bits
    s0, xs0, us0, vs0, lb0 = _solve(capi, pp, mu, theta, explicit_dyn=0)
    assert s0.used_fast_path() and not s0.used_explicit_path()
    if nx != 76:
        assert np.array_equal(xs, xs0) and np.array_equal(us, us0) and np.array_equal(vs, vs0) and np.array_equal(lb, lb0)
        assert np.array_equal(g, s0.download_gains())
    else:   # the general nx = 76 kernel has no 2nd / 3rd order Neumann step (it takes the LDL^T solve there): same
            # numbers to round-off, not the same bits
        for a, b in ((xs, xs0), (us, us0), (lb, lb0)):
            assert np.abs(a - b).max() <= 1e-11 * max(1.0, np.abs(a).max())
    # the promise flag
    s1 = capi.Solver.for_problems(pp, flags=capi.EXPLICIT_DYNAMICS)
    s1.upload(pp); s1.backward(mu, mu)
    xs1, us1, vs1, lb1 = s1.forward(theta)
    assert s1.used_explicit_path() and np.array_equal(xs1, xs) and np.array_equal(lb1, lb)
    # one knot with E != -I: the check must switch the path off after the re-upload
    o = pp.layout.offsets
    pp.knots[B - 1, T // 2, o["E"] + 1] = 1e-3
    s.upload(pp); s.backward(mu, mu)
    xs2, us2, vs2, lb2 = s.forward(theta)
    assert s.used_fast_path() and not s.used_explicit_path()
    _check(oracle, pp, mu, xs2, us2, vs2, lb2, theta)
    for x in (s, s0, s1):
        x.close()
