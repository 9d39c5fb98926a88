"""Generates the committed golden fixtures tests/golden/*.npz.

The reference holds NO golden vectors for the gar solver and cannot be built here (Eigen is
absent), so these fixtures are produced by the repo's own CPU oracle (oracle/, a restatement of
riccati-kernel.hxx / proximal-riccati.hxx) and cross-checked against an independent dense
LAPACK solve of the reference's `lqrDenseMatrix` KKT system (utils.hxx:194-267), which is stored
next to the oracle's answer.  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

CASES = {
    # name: (nx, nu, nc, nth, T, mu, seed, dist)
    "c1_nx36_nu12_T40": (36, 12, 0, 0, 40, 1e-11, 42, "ref"),
    "c2_nx12_nu4_T50": (12, 4, 0, 0, 50, 1e-11, 43, "ref"),
    "c3small_nx36_nu12_nc6_nth4_T20": (36, 12, 6, 4, 20, 1e-11, 44, "ref"),
    "explicit_nx8_nu3_nc2_T15": (8, 3, 2, 0, 15, 1e-8, 45, "explicit"),
}


def load_case(path):
    import garb200_pkg
    garb200_pkg.load()
    from aligator_swk_b200.layout import KnotLayout
    from aligator_swk_b200.synth import PackedProblems
    z = np.load(path)
    nx, nu, nc, nth, T, nc0 = [int(v) for v in z["dims"]]
    pp = PackedProblems(KnotLayout(nx, nu, nc, nth), T, nc0, z["nu_t"].astype(np.int32),
                        z["nc_t"].astype(np.int32), z["knots"], z["G0"], z["g0"])
    theta = z["theta"] if z["theta"].size else None
    return pp, float(z["mu"]), theta, {k: z[k] for k in z.files}


def main():
    import garb200_pkg
    garb200_pkg.load()
    import scipy.linalg as sl
    import pyoracle as po
    from aligator_swk_b200.synth import generate_problems
    po.build()
    for name, (nx, nu, nc, nth, T, mu, seed, dist) in CASES.items():
        pp = generate_problems(1, T, nx, nu, nc, nth, seed=seed, dist=dist)
        rng = np.random.default_rng(seed + 1000)
        theta = rng.standard_normal(nth) if nth else None
        out = po.prox_solve(pp, mu, mu, theta=theta)
        save = dict(dims=np.array([nx, nu, nc, nth, T, pp.nc0]), nu_t=pp.nu_t, nc_t=pp.nc_t,
                    knots=pp.knots, G0=pp.G0, g0=pp.g0, mu=mu,
                    theta=theta if theta is not None else np.zeros(0),
                    xs=out["xs"], us=out["us"], vs=out["vs"], lbdas=out["lbdas"], kkt=out["kkt"])
        if nth == 0:
            mat, rhs = po.dense_kkt(pp, mu, mu)
            xs, us, vs, lb = po.dense_solution_to_traj(pp, sl.solve(mat, -rhs, assume_a="sym"))
            save.update(dense_xs=xs, dense_us=us, dense_vs=vs, dense_lbdas=lb)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **save)
        print(name, "kkt", out["kkt"])


if __name__ == "__main__":
    main()
