#!/usr/bin/env bash
# 1-GPU check: the whole GPU suite, smoke(), then config 2 for 50 rounds (ROC-AUC trajectory of the hyper mode).
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu_n1.log
cat gpurun_out/pytest_gpu_n1.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench/configs.py --config 2 --rounds 50 --warmup 0 2>gpurun_out/cfg2_50_n1.err | tail -1 > gpurun_out/cfg2_r2_50rounds_n1.json
cut -c1-800 gpurun_out/cfg2_r2_50rounds_n1.json
