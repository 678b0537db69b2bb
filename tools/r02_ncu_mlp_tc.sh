#!/bin/bash
# ncu --set full of the tcgen05 MLP kernel on C5 (1 M x 50, cost-only).  Usage: gpurun -- 'bash tools/r02_ncu_mlp_tc.sh [prec]'
set -u
PREC=${1:-tf32}
D=gpurun_out/ncu_mlp_tc
mkdir -p $D
CMD="python bench.py --workload c5 --precision $PREC --steps 3 --warmup 3 --skip-cpu"
$CMD > $D/plain_$PREC.json 2> $D/plain.err && \
ncu --set full --clock-control none --import-source on -k regex:rollout_mlp_tc -s 3 -c 1 -f -o $D/prof_c5_$PREC $CMD > $D/full.log 2>&1
echo "full exit $?"
ls -la $D
