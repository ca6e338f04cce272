#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for m in TransformerModel RNNModel CNNModel; do PROF_S=$([ $m = CNNModel ] && echo 4 || echo 8) python scripts/prof_train.py $m 1280 1 2>&1 | grep -E "plan|launch" | tail -2; done | tee gpurun_out/steps.log
python scripts/prof_train.py CNNModel 640 1 > gpurun_out/prof_cnn_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:train_kernel -s 2 -c 1 -o gpurun_out/prof_cnn -f python scripts/prof_train.py CNNModel 640 1 > gpurun_out/ncu_cnn.log 2>&1
tail -n 2 gpurun_out/ncu_cnn.log
