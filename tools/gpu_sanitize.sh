#!/bin/bash
# compute-sanitizer evidence (run under gpurun): memcheck, racecheck, synccheck on tools/sanitize_cases.py
mkdir -p gpurun_out
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error" gpurun_out/sanitizer_$tool.log | sort | uniq -c | head -12
done
