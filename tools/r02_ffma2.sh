#!/bin/bash
# FFMA2 build: parity (full -m gpu suite incl. the new shape grid), A/B table of the MLP tiles, default bench with the table
set -u
mkdir -p gpurun_out/ffma2
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/ffma2/pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/ffma2/pytest.log
timeout 600 python tools/c2_ab.py > gpurun_out/ffma2/ab.log 2>&1; echo "ab exit $?"; grep -E "peak|tune=\(\) " gpurun_out/ffma2/ab.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/ffma2/bench_default.json 2> gpurun_out/ffma2/bench_default.err; echo "bench exit $?"
python tools/bench_table.py gpurun_out/ffma2/bench_default.json
