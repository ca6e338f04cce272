#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_har_native.py -m gpu -q 2>&1 | tail -12
timeout 300 python bench/har_kernels.py 2>&1 | tail -2 | tee gpurun_out/har_kernels.json
timeout 300 python bench/har_step.py 2>&1 | tail -2 | tee gpurun_out/har_step.json
