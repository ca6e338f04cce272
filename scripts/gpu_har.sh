#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_linear.py tests/test_har_native.py -m gpu -q 2>&1 | tail -8
timeout 300 python bench/har_step.py 2>&1 | tail -1 | tee gpurun_out/har_step.json
ATTACKFL_HAR_FUSE_LN=0 timeout 300 python bench/har_step.py 2>&1 | tail -1
