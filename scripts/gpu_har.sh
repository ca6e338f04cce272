#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_har_native.py tests/test_tc_linear.py -m gpu -q 2>&1 | tail -4
timeout 300 python bench/tc_linear.py 2>&1 | tail -4 | cut -c1-125
timeout 300 python bench/har_kernels.py 2>&1 | tail -1 | tee gpurun_out/har_kernels.json | cut -c330-700
timeout 300 python bench/har_step.py 2>&1 | tail -1 | tee gpurun_out/har_step.json
