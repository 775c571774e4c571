#!/bin/bash
# parity tests + short bench + one full ncu capture of the streaming kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_quick.json; grep -o '"roofline.*' gpurun_out/bench_quick.json | cut -c1-500; tail -5 gpurun_out/bench_quick.err
SHORT="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$SHORT > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fbank_stream -s 3 -c 1 -f -o gpurun_out/prof_stream $SHORT > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"
