#!/bin/bash
set -u
mkdir -p gpurun_out/tg
timeout 900 python -m pytest tests/test_gpu_shapes.py -q -k "two_group or resident" > gpurun_out/tg/pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/tg/pytest.log
timeout 600 python tools/c2_ab.py > gpurun_out/tg/ab.log 2>&1; echo "ab exit $?"; grep -E "peak|N= 1048576|N=   65536|N=   16384" gpurun_out/tg/ab.log
./tools/mb_layer2 > gpurun_out/tg/mb_layer2.log 2>&1; head -5 gpurun_out/tg/mb_layer2.log
