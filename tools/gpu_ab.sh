#!/bin/bash
# A/B on the GPU box: parity tests, then short benches with tuning knobs.
# usage: gpu_ab.sh "<pytest args>" "<workload> [--tune k=v ...]" ...
mkdir -p gpurun_out
PT="$1"; shift
if [ -n "$PT" ]; then
  timeout 1200 python -m pytest $PT -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
fi
i=0
for spec in "$@"; do
  i=$((i+1))
  timeout 300 python bench.py --workload $spec --steps 5 --warmup 3 --no-e2e --cpu-budget 1 --parity-k 0 --others '' > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err; echo "[$i] $spec rc=$?"; tail -2 gpurun_out/ab_$i.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_$i.json")); r=d["roofline"]
    print("   solves/s %.0f ms/step %.3f bwd %.3f fwd %.3f fp64 %.3f hbm %.3f"%(d["value"], d["ms_per_step"], r["kernel_ms"], r["forward_kernel_ms"], r["fp64_frac"], r["hbm_frac"]), d["kkt_max"], d["gpu_launches"])
except Exception as e: print("   ", e)
PY
done
