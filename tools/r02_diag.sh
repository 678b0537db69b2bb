#!/bin/bash
# Round-2 diagnosis visit (verdict next #1): does the FP32 GRU result move under ncu, and which side moves
# (GPU kernel or CPU oracle)?  Then compute-sanitizer over every kernel family.
# usage: gpurun --timeout 2400 -- 'bash tools/r02_diag.sh [repro] [sanitize]'
set -u
what="${*:-repro sanitize}"
mkdir -p gpurun_out/repro gpurun_out/sanitizer
NCU="ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv"
if [[ $what == *repro* ]]; then
  lscpu | head -20 > gpurun_out/repro/lscpu.txt
  timeout 300 python tools/gru_repro.py plain1 > gpurun_out/repro/plain1.log 2>&1; echo "plain1 exit $?"
  timeout 600 $NCU --log-file gpurun_out/repro/launches_ncu1.csv python tools/gru_repro.py ncu1 > gpurun_out/repro/ncu1.log 2>&1; echo "ncu1 exit $?"
  timeout 300 python tools/gru_repro.py plain2 > gpurun_out/repro/plain2.log 2>&1; echo "plain2 exit $?"
  timeout 600 ncu --metrics gpu__time_duration.sum -c 400 --csv --log-file gpurun_out/repro/launches_ncu2.csv python tools/gru_repro.py ncu2 > gpurun_out/repro/ncu2.log 2>&1; echo "ncu2 exit $?"
  for pair in "plain1 plain2" "plain1 ncu1" "plain1 ncu2"; do
    echo "== $pair"; python tools/gru_repro.py --cmp $pair
  done > gpurun_out/repro/cmp.log 2>&1
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/repro/smoke_plain.log 2>&1; echo "smoke plain exit $?"
  timeout 600 $NCU --log-file gpurun_out/repro/launches_smoke.csv python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/repro/smoke_ncu.log 2>&1; echo "smoke ncu exit $?"
  cat gpurun_out/repro/cmp.log
  grep -h "smoke GRU" gpurun_out/repro/smoke_plain.log gpurun_out/repro/smoke_ncu.log
fi
if [[ $what == *sanitize* ]]; then
  for tool in memcheck racecheck initcheck synccheck; do
    for grp in "fp32:mlp2 mlp rnn aux" "tc:rnn_tc mlp_tc lstm_tc"; do
      tag=${grp%%:*}; fams=${grp#*:}
      timeout 900 compute-sanitizer --tool $tool --print-limit 40 --log-file gpurun_out/sanitizer/${tool}_${tag}.log \
          python tools/sanitize_cases.py $fams > gpurun_out/sanitizer/${tool}_${tag}.out 2>&1
      echo "$tool $tag exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitizer/${tool}_${tag}.log | tail -1)"
    done
  done
fi
ls -la gpurun_out/repro gpurun_out/sanitizer
