#!/bin/bash
# ncu --set full (+ source page) of the tcgen05 GRU kernel on C3 tf32x3.  Usage: gpurun -- 'bash tools/r02_ncu_tc.sh [prec]'
set -u
PREC=${1:-tf32x3}
mkdir -p gpurun_out/ncu_tc
CMD="python bench.py --workload c3 --precision $PREC --steps 3 --warmup 3 --skip-cpu"
$CMD > gpurun_out/ncu_tc/plain_$PREC.json 2> gpurun_out/ncu_tc/plain.err && \
ncu --set full --clock-control none --import-source on -k regex:rollout_rnn_tc -s 3 -c 1 -f -o gpurun_out/ncu_tc/prof_c3_$PREC $CMD > gpurun_out/ncu_tc/full.log 2>&1
echo "full exit $?"
ls -la gpurun_out/ncu_tc
