#!/bin/bash
# One full ncu capture of the streaming kernel on the bench command (launch 12 of the run: after the warm-up launches, on the
# rotating 600 MB working set) + the launch list.  Only after the same command has exited 0 without ncu.
mkdir -p gpurun_out
SHORT="python bench.py --steps 6 --warmup 3 --repeats 3 --sustained-seconds 0 --loader-seconds 0 --no-cpu-baseline"
$SHORT > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:fbank_stream -s 12 -c 2 -f -o gpurun_out/prof_stream $SHORT > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out/prof_stream.ncu-rep
