#!/bin/bash
# One gpurun --gpus N call: plain-copy microbenchmark, C5 training arms and bench.py under torchrun on N GPUs of one box.
# usage: bash scripts/gpu_multi.sh N [what...]   (what: h2d c5 bench; default all)
N=$1; shift; WHAT=${@:-h2d c5 bench}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1; nproc >> gpurun_out/topo_n$N.txt
for w in $WHAT; do
  case $w in
    h2d) timeout 300 $TR --master-port 29511 scripts/h2d_microbench.py > gpurun_out/h2d_n$N.json 2> gpurun_out/h2d_n$N.err; echo "h2d rc=$?"; cat gpurun_out/h2d_n$N.json | cut -c1-600;;
    c5) C5_STEPS=20 timeout 900 $TR --master-port 29512 scripts/c5_train.py > gpurun_out/c5_n$N.log 2>&1; echo "c5 rc=$?"; tail -1 gpurun_out/c5_n$N.log | cut -c1-900;;
    bench) timeout 600 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['repeats']['per_rank_ms_per_step_median'], 'e2e', d['e2e']['value'], 'resident', d['e2e']['features_stay_on_gpu']['value'], 'loader', (d['e2e']['dataloader'] or {}).get('value'), 'sustained', d.get('sustained',{}).get('clips_per_s'))";;
  esac
done
