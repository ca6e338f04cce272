#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_har_native.py tests/test_tc_linear.py -m gpu -q 2>&1 | tail -40
