"""One tiny case per kernel variant; no torch import.  Written for compute-sanitizer
   (compute-sanitizer --tool racecheck python tools/sanitize_cases.py) — which is CLOSED on this GPU pool
   (profiles/r02_sanitizer.txt); tests/test_gpu_determinism.py runs the same CASES repeatedly instead."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np
import garb200_pkg
garb200_pkg.load()
from aligator_swk_b200 import capi
from aligator_swk_b200.synth import generate_problems

mu = 1e-9
CASES = [  # nx, nu, nc, nth, T, B, legs, dist, flags, tune
    (12, 4, 0, 0, 4, 3, 0, "ref", 0, {}),                       # sweep_kernel<12,4> + einv_kernel<12> + forward_small
    (36, 12, 0, 0, 3, 2, 0, "ref", 0, {}),                      # sweep_kernel<36,12> + einv_dmma + forward_kernel
    (36, 12, 0, 0, 3, 4, 0, "ref", 0, {"group": 3}),            # lockstep groups
    (36, 12, 6, 4, 3, 2, 0, "ref", 0, {}),                      # EG variant with theta columns
    (36, 12, 0, 0, 3, 2, 0, "explicit_psd", 0, {}),             # explicit-dynamics path, order-2/3 Neumann
    (36, 12, 0, 0, 3, 2, 0, "explicit", 0, {}),                 # indefinite Rhat: Bunch-Kaufman fallback
    (36, 12, 0, 0, 3, 2, 0, "explicit_psd", 0, {"explicit_dyn": 0}),   # LDL^T route of Lambda on the general path
    (76, 32, 0, 0, 2, 1, 0, "ref", 0, {}),                      # sweep_big_kernel + einv_cta_dmma
    (36, 12, 0, 0, 11, 1, 3, "ref", 0, {}),                     # leg sweep + condensed_fast_kernel
    (8, 3, 2, 1, 4, 2, 0, "ref", 0, {}),                        # generic backward_kernel
    (8, 3, 0, 0, 7, 2, 2, "ref", 0, {}),                        # generic legs + condensed_kernel + collapse
]
def main():
    for (nx, nu, nc, nth, T, B, legs, dist, flags, tune) in CASES:
        pp = generate_problems(B, T, nx, nu, nc, nth, seed=3, dist=dist)
        s = capi.Solver.for_problems(pp, num_legs=legs, flags=flags)
        s.set_tuning(**tune)
        s.upload(pp)
        s.backward(mu, mu)
        th = np.ones((B, nth)) if nth else None
        xs, us, vs, lb = s.forward(th)
        k = s.kkt_error(mu, mu, th)
        if legs >= 2:
            s.collapse_feedback()
        g = s.download_gains()
        print(f"nx={nx} nu={nu} nc={nc} nth={nth} T={T} B={B} legs={legs} {dist} {tune}: fast={s.used_fast_path()} "
              f"explicit={s.used_explicit_path()} kkt={k.max():.2e} flags={s.get_flags().max()}", flush=True)
        assert np.isfinite(xs).all() and np.isfinite(g).all()
        s.close()
    # device generator + detection kernel
    s = capi.Solver(12, 4, horizon=3, batch=2)
    s.generate(1, 2); s.backward(mu, mu); s.forward(); print("generate + detect:", s.used_explicit_path(), flush=True); s.close()
    print("sanitize cases done")


if __name__ == "__main__":
    main()
