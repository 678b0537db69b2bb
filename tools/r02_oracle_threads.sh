#!/bin/bash
# How often does the CPU oracle's GRU digest move in smoke(): all host threads against one thread, fresh process each time.
set -u
D=gpurun_out/oracle_threads
mkdir -p $D
for thr in 0 1; do
  n=${1:-14}; [ $thr = 1 ] && n=${2:-$n}; for i in $(seq 1 $n); do
    MPCDL_SMOKE_THREADS=$thr timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep "GRU_64_4_4_16_False: N=" | sed "s/^/threads=$thr run $i: /" >> $D/runs.txt
  done
done
echo "all threads: moved $(grep 'threads=0' $D/runs.txt | grep -vc 'oracle 341e0dacfd') of $(grep -c 'threads=0' $D/runs.txt)"
echo "one thread : moved $(grep 'threads=1' $D/runs.txt | grep -vc 'oracle 341e0dacfd') of $(grep -c 'threads=1' $D/runs.txt)"
grep -v "oracle 341e0dacfd" $D/runs.txt
