#!/bin/bash
# Assemble profiles/<tag>/ from the files a `gpurun -- bash scripts/gpu_prof.sh` (or gpu_round.sh) call left in gpurun_out/.
# usage: bash scripts/make_profile.sh r02a
set -e
TAG=$1; OUT=profiles/$TAG; REP=gpurun_out/prof_stream.ncu-rep
KEY=fbank_stream_kernelILb0ELb1ELb1EfLb0E
mkdir -p $OUT /tmp/mkprof
for f in bench.json bench_ref.json launches.csv lscpu.txt smi.txt sweep.json; do [ -f gpurun_out/$f ] && cp gpurun_out/$f $OUT/ || true; done
ncu -i $REP --page raw --csv > /tmp/mkprof/raw.csv
ncu -i $REP --page details > $OUT/fbank_stream_kernel.details.txt
ncu -i $REP --page source --csv > /tmp/mkprof/src.csv
rm -f $OUT/ncu_summary.json
python scripts/ncu_raw_summary.py /tmp/mkprof/raw.csv fbank_stream_kernel $OUT/ncu_summary.json $TAG
cp $OUT/ncu_summary.json profiles/ncu_summary.json
python scripts/ncu_source_summary.py /tmp/mkprof/src.csv > $OUT/fbank_stream_kernel.source_summary.txt
(cd /tmp/mkprof && rm -f *.cubin && cuobjdump -xelf all $OLDPWD/psla_b200/csrc/libpsla_frontend.so > /dev/null && nvdisasm --print-line-info psla_frontend.sm_100a.cubin > dis.txt)
python scripts/ncu_line_profile.py /tmp/mkprof/src.csv /tmp/mkprof/dis.txt $KEY 60 > $OUT/fbank_stream_kernel.line_profile.txt
ls -la $OUT | tail -8
