#!/bin/bash
# Round evidence run (under gpurun, one GPU): parity tests, smoke, the default bench line, the
# reference arm, the target-shape bench, then the ncu launch lists and --set full captures.
mkdir -p gpurun_out
if [ "$1" != "ncuonly" ]; then
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default rc=$?"; tail -3 gpurun_out/bench_default.err; cat gpurun_out/bench_default.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; cat gpurun_out/bench_reference.json
for wl in c2 c1 c3 c4 c5; do
  timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --cpu-budget 3 --others '' > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?"
done
timeout 600 python bench.py --horizons > gpurun_out/bench_horizons.json 2> gpurun_out/bench_horizons.err; echo "horizons rc=$?"
fi
if [ "$1" == "ncu" ] || [ "$1" == "ncuonly" ]; then
  # name:bench-arguments  (short runs: no e2e, no parity sample, no other_configs -> only the timed kernels launch)
  for spec in "c1b:--workload c1b" "c2:--workload c2" "c3:--workload c3" "c5:--workload c5" "c1bx:--workload c1b --dist explicit_psd" "c2x:--workload c2 --dist explicit_psd"; do
    wl=${spec%%:*}; skip=3; case $wl in *x) skip=2;; esac   # kernels per step: pre-pass + sweep + forward (no pre-pass for E = -I)
    args="${spec#*:} --no-e2e --cpu-budget 1 --parity-k 0 --others ''"
    eval timeout 300 python bench.py $args --steps 2 --warmup 3 > gpurun_out/plain_$wl.log 2>&1 && \
    eval timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py $args --steps 2 --warmup 3 > gpurun_out/ncu_launch_$wl.log 2>&1; echo "launch list $wl rc=$?"
    eval timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sweep_\|einv_\|forward_" -s $skip -c $skip -f -o gpurun_out/full_$wl \
      python bench.py $args --steps 1 --warmup 1 > gpurun_out/ncu_full_$wl.log 2>&1; echo "set full $wl rc=$?"
    # the .ncu-rep files (with source) exceed what gpurun copies back: keep the raw-page CSV (+ the source page of c1b)
    ncu -i gpurun_out/full_$wl.ncu-rep --page raw --csv > gpurun_out/full_$wl.raw.csv 2>/dev/null
    if [ "$wl" == "c1b" ] || [ "$wl" == "c2" ]; then ncu -i gpurun_out/full_$wl.ncu-rep --page source --csv > gpurun_out/full_$wl.source.csv 2>/dev/null; fi
    rm -f gpurun_out/full_$wl.ncu-rep
  done
fi
