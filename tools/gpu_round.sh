#!/bin/bash
# One GPU call: parity tests, short benches, ncu captures of the sweep kernels (run under gpurun).
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
for wl in c2 c1b; do
  timeout 200 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --cpu-budget 2 --parity-k 0 --others '' > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?"; tail -3 gpurun_out/bench_$wl.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$wl.json")); r=d["roofline"]
    print("$wl", "solves/s %.0f ms/step %.2f bwd %.2f fwd %.2f fp64 %.3f hbm %.3f"%(d["value"], d["ms_per_step"], r["kernel_ms"], r["forward_kernel_ms"], r["fp64_frac"], r["hbm_frac"]), d["kkt_max"], d["gpu_launches"])
except Exception as e: print("$wl", e)
PY
done
if [ "$1" == "ncu" ]; then
  timeout 200 python bench.py --workload c1b --batch 296 --steps 1 --warmup 1 --no-e2e --cpu-budget 1 --parity-k 0 --others '' > gpurun_out/plain_c1b.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|einv_kernel" -c 2 -f -o gpurun_out/prof_c1b python bench.py --workload c1b --batch 296 --steps 1 --warmup 1 --no-e2e --cpu-budget 1 --parity-k 0 --others '' > gpurun_out/ncu_c1b.log 2>&1; echo ncu c1b rc=$?
  ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|einv_kernel" -c 2 -f -o gpurun_out/prof_c2 python bench.py --workload c2 --batch 4736 --steps 1 --warmup 1 --no-e2e --cpu-budget 1 --parity-k 0 --others '' > gpurun_out/ncu_c2.log 2>&1; echo ncu c2 rc=$?
fi
