#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_linear.py tests/test_har_native.py -m gpu -q 2>&1 | tail -12
timeout 300 python bench/tc_linear.py 2>&1 | tail -4 | tee gpurun_out/tc_linear.jsonl | cut -c1-420
timeout 300 python bench/har_step.py 2>&1 | tail -1 | tee gpurun_out/har_step.json
