#!/bin/bash
# GPU visit: full -m gpu suite (no -x: every failure listed), then the oracle probe plain and 3x under ncu.
set -u
mkdir -p gpurun_out/probe
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python tools/oracle_probe.py plain > gpurun_out/probe/plain.log 2>&1; echo "plain exit $?"
for i in 1 2 3; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/probe/launches_ncu$i.csv \
      python tools/oracle_probe.py ncu$i > gpurun_out/probe/ncu$i.log 2>&1; echo "ncu$i exit $?"
done
grep -h "rep \|MOVED\|stable\|ops:\|threads" gpurun_out/probe/*.log | cut -c1-330
