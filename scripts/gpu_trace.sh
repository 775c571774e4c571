#!/bin/bash
# parity tests + per-warp timeline of the streaming kernel + short bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 600 python scripts/gpu_trace.py > gpurun_out/trace.log 2>&1; echo "trace rc=$?"; tail -60 gpurun_out/trace.log
timeout 300 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_quick.json; grep -o '"roofline.*' gpurun_out/bench_quick.json | cut -c1-400; tail -5 gpurun_out/bench_quick.err
