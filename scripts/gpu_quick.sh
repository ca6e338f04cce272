#!/usr/bin/env bash
# Quick GPU loop: GPU tests + per-model step timing + short bench.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
for m in TransformerModel RNNModel CNNModel; do python scripts/prof_train.py $m 1280 1 2>&1 | tail -3; done | tee gpurun_out/steps.log
timeout 900 python bench.py --steps ${BENCH_STEPS:-3} --warmup 3 2>&1 | tail -3 | tee gpurun_out/bench.log
