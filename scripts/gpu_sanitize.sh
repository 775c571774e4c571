#!/bin/bash
# compute-sanitizer memcheck + racecheck (+ synccheck) over scripts/sanitize_case.py; logs -> gpurun_out/sanitize_*.log.
# The streaming kernel aliases three shared-memory regions per warp (staged samples -> exchange buffer -> power spectra)
# ordered only by __syncwarp(), and consecutive launches overlap: exactly what racecheck / memcheck are for.
mkdir -p gpurun_out
CS=${CS:-/usr/local/cuda/bin/compute-sanitizer}
timeout 120 python scripts/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
for tool in memcheck racecheck synccheck; do
  timeout 1500 $CS --tool $tool --print-limit 20 --log-file gpurun_out/sanitize_$tool.log python scripts/sanitize_case.py > gpurun_out/sanitize_$tool.out 2>&1
  echo "$tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitize_$tool.log | tail -1)"
done
