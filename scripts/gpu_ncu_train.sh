#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PROF_S=${PROF_S:-8}
python scripts/prof_train.py TransformerModel 1280 1 > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:train_kernel -s 2 -c 1 -o gpurun_out/prof_train3 -f \
    python scripts/prof_train.py TransformerModel 1280 1 > gpurun_out/ncu_train3.log 2>&1
cat gpurun_out/prof_plain.log; tail -n 3 gpurun_out/ncu_train3.log
