#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, timing sweep, ncu launch list + full capture.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh [tests] [bench] [sweep] [ncu]'
set -u
mkdir -p gpurun_out
what="${*:-tests bench sweep ncu}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.csv 2>&1
if [[ $what == *tests* ]]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
fi
if [[ $what == *bench* ]]; then
  timeout 600 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 exit $?"
  timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/bench_c2_ref.json 2>> gpurun_out/bench_c2.err
  for wl in c5 c3 c3l c5g c4; do
    timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl exit $?"
  done
  for prec in tf32x3 tf32; do
    for wl in c2 c5 c3 c3s c3l c5g; do
      timeout 600 python bench.py --workload $wl --precision $prec --steps 5 --warmup 3 --skip-cpu > gpurun_out/bench_${wl}_$prec.json 2> gpurun_out/bench_${wl}_$prec.err; echo "bench $wl $prec exit $?"
    done
  done
fi
if [[ $what == *sweep* ]]; then
  timeout 600 python tools/gpu_quick.py > gpurun_out/sweep.log 2>&1; echo "sweep exit $?"
fi
if [[ $what == *ncu* ]]; then
  timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/plain_c2.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
      python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/ncu_launches_c2.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:rollout_mlp -s 3 -c 2 -f -o gpurun_out/prof_c2 \
      python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/ncu_full_c2.log 2>&1
  timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --skip-cpu > gpurun_out/plain_c3.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:rollout_rnn -s 1 -c 1 -f -o gpurun_out/prof_c3 \
      python bench.py --workload c3 --steps 2 --warmup 3 --skip-cpu > gpurun_out/ncu_full_c3.log 2>&1
fi
if [[ $what == *tcprof* ]]; then
  timeout 300 python bench.py --workload c5 --precision tf32x3 --steps 2 --warmup 3 --skip-cpu > gpurun_out/plain_c5_tf32x3.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:rollout_mlp_tc -s 1 -c 1 -f -o gpurun_out/prof_c5_tf32x3 \
      python bench.py --workload c5 --precision tf32x3 --steps 2 --warmup 3 --skip-cpu > gpurun_out/ncu_full_c5_tf32x3.log 2>&1
  timeout 300 python bench.py --workload c3 --precision tf32x3 --steps 2 --warmup 3 --skip-cpu > gpurun_out/plain_c3_tf32x3.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:rollout_rnn_tc -s 1 -c 1 -f -o gpurun_out/prof_c3_tf32x3 \
      python bench.py --workload c3 --precision tf32x3 --steps 2 --warmup 3 --skip-cpu > gpurun_out/ncu_full_c3_tf32x3.log 2>&1
fi
ls -la gpurun_out
if [[ $what == *c4bf16* ]]; then
  timeout 900 python -m pytest tests/test_gpu_lstm256.py -m gpu -x -q > gpurun_out/pytest_lstm256.log 2>&1; echo "pytest lstm256 exit $?" | tee -a gpurun_out/pytest_lstm256.log
  timeout 600 python bench.py --workload c4 --precision bf16 --steps 10 --warmup 3 > gpurun_out/bench_c4_bf16.json 2> gpurun_out/bench_c4_bf16.err; echo "bench c4 bf16 exit $?"
  timeout 300 python bench.py --workload c4 --precision bf16 --steps 2 --warmup 3 --skip-cpu > gpurun_out/plain_c4_bf16.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/launches_c4_bf16.csv \
      python bench.py --workload c4 --precision bf16 --steps 2 --warmup 3 --skip-cpu > gpurun_out/ncu_launches_c4_bf16.log 2>&1
  timeout 120 python tools/tc_prof.py LSTM_256_4_4_0_False bf16 18944 10 > gpurun_out/plain_prof_c4_bf16.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:rollout_lstm_tc -s 1 -c 1 -f -o gpurun_out/prof_c4_bf16 \
      python tools/tc_prof.py LSTM_256_4_4_0_False bf16 18944 10 > gpurun_out/ncu_full_c4_bf16.log 2>&1
fi
