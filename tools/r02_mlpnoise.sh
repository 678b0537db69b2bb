#!/bin/bash
set -u
mkdir -p gpurun_out/mlpnoise
timeout 900 python -m pytest tests -m gpu -q -x -k "tensor_core or tc or tf32 or t3 or precision" > gpurun_out/mlpnoise/pytest.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/mlpnoise/pytest.log
