"""Per-phase cycle breakdown of the specialised sweep (diagnostic; run on the GPU box):
   python tools/phase_timers.py c1b 296 [group=1 ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, garb200_pkg
garb200_pkg.load()
from aligator_swk_b200 import capi
w = dict(bench.WORKLOADS[sys.argv[1]]); B = int(sys.argv[2]) if len(sys.argv) > 2 else w["batch"]
s = capi.Solver(w["nx"], w["nu"], w["nc"], w["nth"], horizon=w["T"], batch=B, flags=capi.PHASE_TIMERS)
kv = {a.split("=")[0]: int(a.split("=")[1]) for a in sys.argv[3:]}
dist = kv.pop("dist", 0)
s.set_tuning(**kv)
s.generate(42, dist)
for _ in range(2):
    s.backward(bench.MU, bench.MU)
s.synchronize_stream_only(); s.phase_cycles()
s.backward(bench.MU, bench.MU); s.synchronize_stream_only()
c = s.phase_cycles().astype(float) / (B * w["T"])
names = ["load", "g1 EinvT P", "g2 Ptilde", "gemv+Lambda", "ldlt", "mirror", "g5 AtV/BtV", "g6 Qhat/Rhat+rhs", "bk factor",
         "bk solve", "L/Apre", "g10 Afb", "g11 Pmat", "store", "init-pre", "initial"]
tot = c[:14].sum()
for n, v in zip(names, c):
    print(f"{n:18s} {v:10.0f} cycles/knot  {100*v/tot:5.1f}%")
print("total per knot", tot, "cycles; initial stage per problem", c[15] * w["T"])
