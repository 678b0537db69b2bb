#!/bin/bash
set -u
mkdir -p gpurun_out/win8
timeout 900 python -m pytest tests -m gpu -q -x -k "tensor_core or tc or tf32 or t3 or precision" > gpurun_out/win8/pytest.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/win8/pytest.log
