#!/bin/bash
# DRAM / L2 traffic of one launch of the streaming kernel per output layout (run only after layout_bench.py exits 0)
mkdir -p gpurun_out
timeout 300 python scripts/layout_bench.py > gpurun_out/layout_bench.log 2>&1 || { echo "layout_bench failed"; tail -5 gpurun_out/layout_bench.log; exit 1; }
tail -1 gpurun_out/layout_bench.log
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum,lts__t_requests_srcunit_tex_op_write.sum,smsp__inst_executed.sum
for L in 0 1 2; do
  timeout 600 ncu --metrics $M --clock-control none --kernel-name-base mangled -k regex:"fbank_stream_kernel.*Li${L}EE" --launch-skip 25 --launch-count 1 --csv \
      --log-file gpurun_out/layout_ncu_${L}.csv python scripts/layout_bench.py > gpurun_out/layout_ncu_${L}.log 2>&1
  echo "ncu layout $L rc=$?"
done
