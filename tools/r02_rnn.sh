#!/bin/bash
set -u
mkdir -p gpurun_out/rnn
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/rnn/pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/rnn/pytest.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/rnn/bench_default.json 2> gpurun_out/rnn/bench_default.err; echo "bench exit $?"
python tools/bench_table.py gpurun_out/rnn/bench_default.json
timeout 300 python bench.py --workload c3l --steps 5 --warmup 3 --skip-cpu > gpurun_out/rnn/bench_c3l.json 2>/dev/null; python tools/bench_table.py gpurun_out/rnn/bench_c3l.json | head -2
