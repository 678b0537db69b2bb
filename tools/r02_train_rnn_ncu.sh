#!/bin/bash
set -u
D=gpurun_out/train_rnn_ncu
mkdir -p $D
CMD="python tools/train_rnn_ncu.py gru 64 1 40"
$CMD > $D/plain.log 2>&1; cat $D/plain.log | tail -1
ncu --set full --clock-control none --import-source on -k regex:train_rnn -s 1 -c 1 -f -o $D/prof_gru64 $CMD > $D/full.log 2>&1; echo "ncu exit $?"
ls -la $D
