#!/bin/bash
# Quick ncu look at one kernel of one workload (under gpurun): time, instructions, issue, smem pipe, bank
# conflicts, DMMA pipe.  usage: gpu_kernel_metrics.sh <kernel regex> <bench args...>
mkdir -p gpurun_out
K="$1"; shift
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed,sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,dram__bytes_read.sum,dram__bytes_write.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:$K -s 1 -c 1 --csv --log-file gpurun_out/kmetrics.csv \
  python bench.py "$@" --steps 1 --warmup 1 --no-e2e --cpu-budget 1 --parity-k 0 --others "" > gpurun_out/kmetrics.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(l for l in open("gpurun_out/kmetrics.csv") if not l.startswith("==")))
h=rows[0]; ik,im,iv,iu=h.index("Kernel Name"),h.index("Metric Name"),h.index("Metric Value"),h.index("Metric Unit")
for r in rows[1:]:
    if len(r)>iv: print("%-40s %-80s %s %s"%(r[ik][:40], r[im][-80:], r[iv], r[iu]))
PY
