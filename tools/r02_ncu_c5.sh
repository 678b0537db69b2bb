#!/bin/bash
# ncu evidence for the default bench workload (C5: rollout_mlp2_kernel<64,8,8,1>): launch list of the bench command, then
# one --set full capture of the kernel (source page included).  Usage: gpurun -- 'bash tools/r02_ncu_c5.sh'
set -u
mkdir -p gpurun_out/ncu_c5
CMD="python bench.py --steps 3 --warmup 3 --no-table --skip-cpu"
$CMD > gpurun_out/ncu_c5/plain.json 2> gpurun_out/ncu_c5/plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_c5/launches.csv $CMD > gpurun_out/ncu_c5/under_ncu.log 2>&1
echo "launch list exit $?"
$CMD > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:rollout_mlp2 -s 3 -c 1 -f -o gpurun_out/ncu_c5/prof_c5 $CMD > gpurun_out/ncu_c5/full.log 2>&1
echo "full exit $?"
python tools/bench_table.py gpurun_out/ncu_c5/plain.json 2>/dev/null | head -5
ls -la gpurun_out/ncu_c5
