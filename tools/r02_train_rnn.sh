#!/bin/bash
set -u
mkdir -p gpurun_out/train_rnn
timeout 900 python -m pytest tests/test_gpu_train_rnn.py tests/test_gpu_train.py -q > gpurun_out/train_rnn/pytest.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/train_rnn/pytest.log
timeout 600 python tools/train_rnn_time.py > gpurun_out/train_rnn/timing.log 2>&1; echo "timing exit $?"; cat gpurun_out/train_rnn/timing.log | tail -20
