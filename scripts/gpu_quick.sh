#!/bin/bash
# quick GPU check: diagnostics + parity tests + short bench (no ncu)
mkdir -p gpurun_out
timeout 600 python scripts/gpu_check.py > gpurun_out/gpu_check.log 2>&1; echo "gpu_check rc=$?"; tail -25 gpurun_out/gpu_check.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cat gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
