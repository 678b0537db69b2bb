#!/bin/bash
# Round-2 GPU visit: parity tests, smoke (plain + under ncu), default bench with the per-workload table, reference arm.
# usage: gpurun --timeout 1800 -- 'bash tools/r02_round.sh [tests] [smoke] [bench] [ref]'
set -u
what="${*:-tests smoke bench ref}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.csv 2>&1
if [[ $what == *tests* ]]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  tail -15 gpurun_out/pytest_gpu.log
fi
if [[ $what == *smoke* ]]; then
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_smoke.csv \
      python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_ncu.log 2>&1; echo "smoke under ncu exit $?"
  diff gpurun_out/smoke.log gpurun_out/smoke_ncu.log > gpurun_out/smoke_diff.txt; cat gpurun_out/smoke.log
fi
if [[ $what == *bench* ]]; then
  timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"
  tail -c 600 gpurun_out/bench_default.err
  python tools/bench_table.py gpurun_out/bench_default.json
fi
if [[ $what == *ref* ]]; then
  timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
  cat gpurun_out/bench_ref.json | head -c 900
fi
