import sys, os
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle")]
import numpy as np, garb200_pkg
garb200_pkg.load()
from aligator_swk_b200 import capi
from aligator_swk_b200.synth import generate_problems
import pyoracle as po
rel = lambda a, r: float(np.abs(a - r).max() / max(np.abs(r).max(), 1e-300))
for (nx, nu, T, mu) in [(36, 12, 20, 1e-9), (36, 12, 100, 1e-9), (36, 12, 100, 1e-11), (36, 12, 100, 1e-6), (12, 4, 50, 1e-9)]:
    pp = generate_problems(2, T, nx, nu, seed=61, dist="explicit")
    for ex in (0, 1):
        for fg in (0, capi.FORCE_GENERIC):
            if fg and ex: continue
            s = capi.Solver.for_problems(pp, flags=fg)
            s.set_tuning(explicit_dyn=ex)
            s.upload(pp); s.backward(mu, mu)
            xs, us, vs, lb = s.forward()
            kkt = s.kkt_error(mu, mu)
            errs = []
            for b in range(2):
                ref = po.prox_solve(pp, mu, mu, b=b)
                errs.append(max(rel(xs[b], ref["xs"]), rel(us[b], ref["us"]), rel(lb[b], ref["lbdas"])))
            print(f"nx={nx} T={T} mu={mu:g} explicit_dyn={ex} generic={bool(fg)} used_ex={s.used_explicit_path()} err {max(errs):.2e} kkt {kkt.max(axis=0)}", flush=True)
            s.close()
