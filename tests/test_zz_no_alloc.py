"""garb200_kkt_error_noalloc (include/garb200.h): the allocation-free variant of the KKT metric
(no counterpart in the reference).  Through the C ABI, against
the existing entry point garb200_kkt_error on the same device data, two calls in a row — the static
device scratch is shared by every call and keeps the last call's contents.

Collected LAST on purpose (file name): a failure here must not hide
the parity tests.  Passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

from aligator_swk_b200.synth import generate_problems

pytestmark = pytest.mark.gpu

RTOL = 1e-10   # the parity tests' tolerance; the two entry points run the same kernel, so they agree to the bit


@pytest.fixture(scope="module")
def capi():
    from aligator_swk_b200 import capi as c
    c.lib()
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return c


def _solved(capi, nx, nu, nc, nth, T, B, mu, seed, theta=None, **gen):
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=seed, **gen)
    s = capi.Solver.for_problems(pp)
    s.upload(pp)
    s.backward(mu, mu)
    s.forward(theta)                       # leaves the trajectories in the solver-owned DEVICE buffers
    return pp, s


def _both(capi, pp, s, mu, theta=None):
    import torch
    ref = s.kkt_error(mu, mu, theta)       # existing entry point (handle path, cudaMalloc'd scratch)
    d_knots, d_G0, d_g0 = s.device_problem()
    d_xs, d_us, d_vs, d_lb = s.device_trajectory()
    th_t = None
    if theta is not None:                  # the caller's own device buffer for theta (an INPUT)
        th_t = torch.as_tensor(np.array(np.broadcast_to(theta, (s.batch, s.nth))), device="cuda")
        torch.cuda.synchronize()
    ragged = not (np.all(pp.nu_t[:-1] == s.nu) and np.all(pp.nc_t[:-1] == s.nc))
    got = capi.kkt_error_noalloc(
        s.nx, s.nu, s.nc, s.nth, s.horizon, s.nc0, s.batch, d_knots, d_G0, d_g0, d_xs, d_us, d_vs, d_lb, mu, mu,
        d_theta=None if th_t is None else th_t.data_ptr(),
        nu_t=pp.nu_t if ragged else None, nc_t=pp.nc_t if ragged else None,
        nu_term=int(pp.nu_t[-1]), nc_term=int(pp.nc_t[-1]))
    return got, ref


def test_noalloc_matches_existing_entry_point_twice_in_a_row(capi):
    mu = 1e-9
    # call 1: constrained + parametric shape, larger batch and horizon (fills more of the scratch)
    theta = np.random.default_rng(0).standard_normal((24, 4))
    pp1, s1 = _solved(capi, 36, 12, 6, 4, 30, 24, mu, seed=3, theta=theta, dist="explicit")
    got1, ref1 = _both(capi, pp1, s1, mu, theta)
    assert np.isfinite(got1).all() and ref1.max() > 0.0
    assert np.allclose(got1, ref1, rtol=RTOL, atol=0.0), np.abs(got1 - ref1).max()
    # call 2, right after: another shape, SMALLER batch and horizon — stale contents of the shared
    # scratch from call 1 (dims arrays, result staging) must not leak into the result
    pp2, s2 = _solved(capi, 12, 4, 0, 0, 7, 5, mu, seed=4)
    got2, ref2 = _both(capi, pp2, s2, mu)
    assert got2.shape == (5, 3)
    assert np.allclose(got2, ref2, rtol=RTOL, atol=0.0), np.abs(got2 - ref2).max()
    # and call 1's data once more: same answer as the first time
    got1b, _ = _both(capi, pp1, s1, mu, theta)
    assert np.array_equal(got1b, got1)
    s1.close(); s2.close()


def test_noalloc_ragged_knots_through_host_dims(capi):
    """per-knot nu_t / nc_t given by the caller (host arrays) instead of (nu, nc, nu_term, nc_term)"""
    from test_oracle import short_horz_pb
    pp = short_horz_pb().pack()
    s = capi.Solver.for_problems(pp)
    s.upload(pp)
    s.backward(1e-12, 1e-12)
    s.forward()
    got, ref = _both(capi, pp, s, 1e-12)
    assert np.allclose(got, ref, rtol=RTOL, atol=0.0)
    s.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(capi):
    """8 (T+1) bytes (256-rounded) + 24 batch bytes must fit GARB200_NOALLOC_SCRATCH_BYTES."""
    pp, s = _solved(capi, 4, 2, 0, 0, 3, 2, 1e-9, seed=1)
    d_knots, d_G0, d_g0 = s.device_problem()
    d_xs, d_us, d_vs, d_lb = s.device_trajectory()
    too_many = capi.NOALLOC_SCRATCH_BYTES // 24 + 1
    with pytest.raises(capi.GarB200Error) as e:       # rejected before any pointer is touched
        capi.kkt_error_noalloc(4, 2, 0, 0, 3, 4, too_many, d_knots, d_G0, d_g0, d_xs, d_us, d_vs, d_lb, 1e-9, 1e-9)
    assert e.value.code == capi.ERR_INVALID and "GARB200_NOALLOC_SCRATCH_BYTES" in str(e.value)
    with pytest.raises(capi.GarB200Error) as e:
        capi.kkt_error_noalloc(4, 2, 0, 0, capi.NOALLOC_SCRATCH_BYTES // 8, 4, 2, d_knots, d_G0, d_g0, d_xs, d_us,
                               d_vs, d_lb, 1e-9, 1e-9)
    assert e.value.code == capi.ERR_INVALID
    # the largest batch that fits is accepted by the size check (it then needs real data: not run here)
    got, ref = _both(capi, pp, s, 1e-9)
    assert np.allclose(got, ref, rtol=RTOL, atol=0.0)
    s.close()
