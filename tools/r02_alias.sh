#!/bin/bash
set -u
mkdir -p gpurun_out/alias
timeout 900 python -m pytest tests -m gpu -q -x -k "tensor_core or tc or noise or tf32 or t3 or precision" > gpurun_out/alias/pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/alias/pytest.log
for prec in tf32x3 tf32; do
  timeout 300 python bench.py --workload c3 --precision $prec --steps 5 --warmup 3 --skip-cpu > gpurun_out/alias/c3_$prec.json 2>/dev/null; python tools/bench_table.py gpurun_out/alias/c3_$prec.json | head -2
done
