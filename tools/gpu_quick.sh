#!/bin/bash
# quick A/B on the GPU box: parity subset + short benches.  usage: gpu_quick.sh [workloads...]
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for wl in "${@:-c2 c1b}"; do for w in $wl; do
  timeout 200 python bench.py --workload $w --steps 5 --warmup 3 --no-e2e --cpu-budget 1 --parity-k 0 --others '' > gpurun_out/q_$w.json 2> gpurun_out/q_$w.err; echo "$w rc=$?"; tail -3 gpurun_out/q_$w.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/q_$w.json")); r=d["roofline"]
    print("$w", "solves/s %.0f ms/step %.2f bwd %.2f fwd %.2f fp64 %.3f hbm %.3f"%(d["value"], d["ms_per_step"], r["kernel_ms"], r["forward_kernel_ms"], r["fp64_frac"], r["hbm_frac"]), d["kkt_max"], d["gpu_launches"])
except Exception as e: print("$w", e)
PY
done; done
