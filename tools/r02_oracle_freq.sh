#!/bin/bash
# How often does the CPU oracle's GRU evaluation move in smoke() under ncu, and does the host thread configuration matter?
# usage: gpurun --timeout 1800 -- 'bash tools/r02_oracle_freq.sh'
set -u
mkdir -p gpurun_out/ofreq
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ofreq/plain.log 2>&1; echo "plain exit $?"
run() {   # tag, env assignments...
  tag=$1; shift
  for i in 1 2 3 4 5 6; do
    env "$@" timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ofreq/launches_${tag}_$i.csv \
        python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ofreq/${tag}_$i.log 2>&1
    echo "$tag $i exit $? : $(grep -c 'ORACLE MOVED' gpurun_out/ofreq/${tag}_$i.log) moved; $(grep 'smoke GRU_64_4_4_16_False:' gpurun_out/ofreq/${tag}_$i.log | head -1 | cut -c1-200)"
  done
}
run default DUMMY=1
run omp1 OMP_NUM_THREADS=1 MKL_NUM_THREADS=1
run nodyn MKL_DYNAMIC=FALSE OMP_DYNAMIC=FALSE
rm -f gpurun_out/ofreq/launches_*.csv
grep -h "ORACLE MOVED" gpurun_out/ofreq/*.log | cut -c1-300
