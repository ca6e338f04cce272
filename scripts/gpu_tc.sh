#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_linear.py -m gpu -q -x 2>&1 | tail -25
