#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (both arms, the driver's command line), sweeps, ncu launch list + one full
# capture of the streaming kernel.  Then: bash scripts/make_profile.sh <tag>
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
lscpu | head -20 > gpurun_out/lscpu.txt; nproc >> gpurun_out/lscpu.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/bench.json
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-300 gpurun_out/bench_ref.json
python scripts/sweep.py > gpurun_out/sweep.log 2>&1; echo "sweep rc=$?"
python scripts/layout_bench.py > gpurun_out/layout_bench.log 2>&1; echo "layout rc=$?"; tail -3 gpurun_out/layout_bench.log
python scripts/parity_report.py gpurun_out/parity.json > gpurun_out/parity.log 2>&1; echo "parity rc=$?"
bash scripts/gpu_prof.sh
ls -la gpurun_out
