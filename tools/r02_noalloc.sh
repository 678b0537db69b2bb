#!/bin/bash
set -u
mkdir -p gpurun_out/noalloc
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/noalloc/alone.log 2>&1; echo "alone exit $?"; tail -15 gpurun_out/noalloc/alone.log
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/noalloc/all.log 2>&1; echo "all exit $?"; tail -5 gpurun_out/noalloc/all.log
