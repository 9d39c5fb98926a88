#!/bin/bash
# Multi-GPU A/B (run under gpurun --gpus N): gather modes, strong scaling, multi-device handle test.
# usage: gpu_multi.sh N ["bench args" ...]   (default: the full list of modes)
N=${1:-2}; shift
mkdir -p gpurun_out
Q="--steps 10 --warmup 3 --no-e2e --parity-k 0 --cpu-budget 1"
i=0
if [ $# -eq 0 ]; then set -- "" "--gather p2p" "--scaling strong" "--workload c2" "--workload c2 --gather p2p" "--workload c2 --scaling strong" "--workload c5 --scaling strong"; fi
for spec in "$@"; do
  i=$((i+1))
  timeout 300 python bench.py --gpus $N $Q --others "" $spec > gpurun_out/multi_${N}_$i.json 2> gpurun_out/multi_${N}_$i.err; echo "[$i] N=$N $spec rc=$?"
  python - "$N" "$i" <<'PY'
import json, sys
try:
    txt = open(f"gpurun_out/multi_{sys.argv[1]}_{sys.argv[2]}.json").read()
    d = json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r = d["roofline"]
    print("    %.0f solves/s  %.3f ms/step  backward kernel %.3f ms  gather=%s scaling=%s global_batch=%s" % (
        d["value"], d["ms_per_step"], r["kernel_ms"], d.get("gather"), d["scaling"], d["config"]["global_batch"]))
except Exception as e:
    print("    ", e)
PY
done
timeout 300 python -m pytest tests/test_gpu_consumers.py -m gpu -q -k multi 2>&1 | tail -2
