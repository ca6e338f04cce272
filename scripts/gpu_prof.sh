#!/usr/bin/env bash
# Full GPU test pass, then an ncu capture of the training kernel on a short workload.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -25 | tee gpurun_out/pytest_gpu.log
echo "== prof plain"
python scripts/prof_train.py TransformerModel 1280 1 > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:train_kernel -s 2 -c 1 -o gpurun_out/prof_train -f \
    python scripts/prof_train.py TransformerModel 1280 1 > gpurun_out/ncu_train.log 2>&1
cat gpurun_out/prof_plain.log
tail -5 gpurun_out/ncu_train.log
for m in RNNModel CNNModel; do python scripts/prof_train.py $m 1280 1 2>&1 | tail -3; done
