"""Developer probe run on the GPU box: prints parity numbers per case without asserting."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import numpy as np
import garb200_pkg; garb200_pkg.load()
from aligator_swk_b200 import capi
from aligator_swk_b200.synth import generate_problems
import pyoracle as po

def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)) if a.size else 0.0

cases = [(2,2,0,0,0,2,1e-13), (2,2,0,0,4,2,1e-12), (12,4,0,0,50,8,1e-11), (36,12,0,0,100,4,1e-11),
         (8,3,2,0,10,2,1e-8), (7,5,3,2,9,2,1e-6), (36,12,6,4,20,2,1e-11), (48,16,0,0,8,2,1e-11)]
if len(sys.argv) > 1:
    cases = cases[: int(sys.argv[1])]
for (nx,nu,nc,nth,T,B,mu) in cases:
    pp = generate_problems(B,T,nx,nu,nc,nth,seed=5)
    theta = np.random.default_rng(1).standard_normal((B,nth)) if nth else None
    t0 = time.time()
    try:
        s = capi.Solver.for_problems(pp, flags=capi.STORE_ALL_VALUES)
        s.upload(pp); s.backward(mu,mu)
        xs,us,vs,lb = s.forward(theta)
    except Exception as e:
        print("CASE", (nx,nu,nc,nth,T,B), "FAILED:", e, "flags", s.get_flags() if 's' in dir() else None); continue
    kk = s.kkt_error(mu,mu,theta)
    errs = []
    for b in range(B):
        ref = po.prox_solve(pp,mu,mu,theta=None if theta is None else theta[b], b=b)
        errs.append([rel(xs[b],ref["xs"]), rel(us[b],ref["us"]), rel(vs[b],ref["vs"]), rel(lb[b],ref["lbdas"])])
        if b == 0:
            g = s.download_gains(0,1)[0]
            ge = [rel(g[t], ref["gains"][t]) for t in range(T+1)]
            v0 = s.get_value(0, 0)
    print("CASE", (nx,nu,nc,nth,T,B,mu), "rel[x,u,v,l] max", np.max(errs,axis=0), "gains rel first/last/max",
          ge[0], ge[-1], max(ge), "kkt gpu", kk.max(axis=0), "oracle kkt", ref["kkt"], "flags", s.get_flags().max(),
          "t=%.2fs" % (time.time()-t0), flush=True)
    s.close()
for kind in (0, 1):
    print("fp64 peak kind", kind, capi.fp64_peak(0, kind), "TFLOP/s", flush=True)
