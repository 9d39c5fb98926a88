"""A/B of the nx = 76 pre-pass variants on the same generated problems (run on the GPU box):
   python tools/dbg_einv.py  -> max relative difference of the trajectories, flags, and vs the CPU oracle on 2 problems"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import garb200_pkg
garb200_pkg.load()
from aligator_swk_b200 import capi
from aligator_swk_b200.synth import generate_problems
import pyoracle as po
mu = 1e-11
pp = generate_problems(3, 12, 76, 32, seed=5)
out = {}
for mode in (0, 3):
    s = capi.Solver.for_problems(pp)
    s.set_tuning(einv_mode=mode)
    s.upload(pp); s.backward(mu, mu)
    out[mode] = s.forward()
    print("mode", mode, "flags", s.get_flags(), "kkt", s.kkt_error(mu, mu).max(axis=0))
    s.close()
for a, b in zip(out[0], out[3]):
    print("rel diff", np.abs(a - b).max() / max(np.abs(a).max(), 1e-300))
for b in range(2):
    ref = po.prox_solve(pp, mu, mu, b=b)
    for mode in (0, 3):
        print("problem", b, "mode", mode, "vs oracle xs", np.abs(out[mode][0][b] - ref["xs"]).max() / np.abs(ref["xs"]).max())
