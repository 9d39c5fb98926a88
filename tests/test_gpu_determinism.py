"""Stand-in for compute-sanitizer racecheck, which is closed on this GPU pool (profiles/r02_sanitizer.txt):
every kernel variant (tools/sanitize_cases.py CASES) solved 8 times on the same bytes must give
bit-identical gains, trajectories and flags.  A shared-memory race, a missing barrier or a read of
uninitialised scratch in the warp-synchronous sweeps shows up as run-to-run differences."""
import os
import sys

import numpy as np
import pytest

from aligator_swk_b200.synth import generate_problems

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from sanitize_cases import CASES  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", CASES, ids=lambda c: "nx%d_nu%d_nc%d_nth%d_T%d_B%d_J%d_%s" % c[:8] + "".join(f"_{k}{v}" for k, v in c[9].items()))
def test_repeated_solves_are_bit_identical(case):
    from aligator_swk_b200 import capi
    nx, nu, nc, nth, T, B, legs, dist, flags, tune = case
    B = max(B, 5)                       # several CTAs in flight, a batch that is not a multiple of the lockstep group
    T = max(T, 6)
    mu = 1e-9
    pp = generate_problems(B, T, nx, nu, nc, nth, seed=9, dist=dist)
    th = np.linspace(-1, 1, B * nth).reshape(B, nth) if nth else None
    s = capi.Solver.for_problems(pp, num_legs=legs, flags=flags)
    s.set_tuning(**tune)
    s.upload(pp)
    first = None
    for rep in range(8):
        s.backward(mu, mu)
        out = [a.copy() for a in s.forward(th)] + [s.download_gains(), s.get_flags()]
        if first is None:
            first = out
            assert all(np.isfinite(a).all() for a in out[:5])
        else:
            for a, b in zip(first, out):
                assert np.array_equal(a, b), f"run {rep} differs from run 0"
    s.close()
