"""Off-grid shapes: backward time per knot through the padded (specialised) path, the same shape on the generic kernel
(GARB200_NO_PADDING) where that exists, and the covering native shape -- VERDICT r1 next #6 ("within 1.5x of the nearest
native shape's per-knot time").  Run on the GPU box: python tools/padded_shapes.py > gpurun_out/padded_shapes.json"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import garb200_pkg
garb200_pkg.load()
from aligator_swk_b200 import capi

MU = 1e-9
def time_shape(nx, nu, nc, nth, T, B, flags=0, reps=5):
    try:
        s = capi.Solver(nx, nu, nc, nth, horizon=T, batch=B, flags=flags)
    except capi.GarB200Error as e:
        return None
    s.generate(7, 0)
    for _ in range(2):
        s.backward(MU, MU); s.forward()
    ts = []
    for _ in range(reps):
        s.backward(MU, MU); s.forward(); s.synchronize_stream_only()
        ts.append(s.last_timings()[0])
    out = {"backward_ms": float(np.median(ts)), "us_per_knot_per_problem": float(np.median(ts)) * 1e3 / (B * T),
           "fast": bool(s.used_fast_path()), "padded_into": s.padded_shape()}
    s.close()
    return out

rows = []
for shape, B, T in [((14, 7, 0, 0), 4096, 50), ((10, 3, 0, 0), 8192, 50), ((30, 10, 0, 0), 2048, 50), ((36, 12, 6, 2), 1024, 50),
                    ((36, 12, 12, 0), 1024, 50), ((20, 8, 3, 1), 1024, 50), ((50, 20, 0, 0), 592, 50), ((64, 24, 0, 0), 592, 50)]:
    pad = time_shape(*shape, T, B)
    gen = time_shape(*shape, T, B, flags=capi.NO_PADDING)
    nat = time_shape(*pad["padded_into"], T, B) if pad and pad["padded_into"] else None
    rows.append({"shape": shape, "batch": B, "horizon": T, "padded": pad, "generic": gen, "native_cover": nat,
                 "padded_vs_native": pad["backward_ms"] / nat["backward_ms"] if pad and nat else None,
                 "generic_vs_padded": gen["backward_ms"] / pad["backward_ms"] if pad and gen else None})
print(json.dumps({"metric": "backward ms (E^-1 pre-pass + sweep), device-resident, median of 5", "mu": MU, "rows": rows}))
