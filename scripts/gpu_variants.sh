#!/bin/bash
# One gpurun call: time every build in psla_b200/csrc/variants/ named on the command line (scripts/variant_bench.py).
# usage: bash scripts/gpu_variants.sh base nowait x_noshift ...
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
: > gpurun_out/variants.jsonl
for v in "$@"; do
  PSLA_FRONTEND_LIB=$PWD/psla_b200/csrc/variants/libpsla_$v.so timeout 300 python scripts/variant_bench.py $v $VB_ARGS >> gpurun_out/variants.jsonl 2> gpurun_out/variant_$v.err || echo "{\"tag\": \"$v\", \"failed\": true}" >> gpurun_out/variants.jsonl
done
cat gpurun_out/variants.jsonl
