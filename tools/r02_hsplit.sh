#!/bin/bash
set -u
D=gpurun_out/${1:-hsplit}
mkdir -p $D
timeout 900 python -m pytest tests -m gpu -q -x -k "tensor_core or tc or noise or tf32 or t3 or precision or no_alloc or static" > $D/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $D/pytest.log
for prec in tf32x3 tf32; do
  for wl in c3 c3l c3s; do
    timeout 300 python bench.py --workload $wl --precision $prec --steps 5 --warmup 3 --skip-cpu > $D/${wl}_$prec.json 2>/dev/null; python tools/bench_table.py $D/${wl}_$prec.json | head -2
  done
done
