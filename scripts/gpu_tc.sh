#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_linear.py -m gpu -q 2>&1 | tail -12
python bench/tc_linear.py 2>&1 | tee gpurun_out/tc_linear.jsonl | tail -6
