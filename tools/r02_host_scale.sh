#!/bin/bash
# usage: gpurun [--gpus N] -- 'bash tools/r02_host_scale.sh N'
set -u
N=${1:-1}
mkdir -p gpurun_out/scale
if [ "$N" = "1" ]; then
  timeout 300 python tools/host_overhead.py > gpurun_out/scale/host_overhead.log 2>&1; cat gpurun_out/scale/host_overhead.log
  timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --skip-cpu > gpurun_out/scale/bench_c2_n1.json 2> gpurun_out/scale/bench_c2_n1.err; echo "c2 exit $?"
  python tools/bench_table.py gpurun_out/scale/bench_c2_n1.json
else
  for wl in ${2:-c5 c2}; do
    timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --steps 20 --warmup 5 \
        > gpurun_out/scale/bench_${wl}_n$N.json 2> gpurun_out/scale/bench_${wl}_n$N.err; echo "$wl n=$N exit $?"
    tail -3 gpurun_out/scale/bench_${wl}_n$N.err | cut -c1-300
    python tools/bench_table.py gpurun_out/scale/bench_${wl}_n$N.json
  done
fi
