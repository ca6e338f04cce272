#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_har_native.py -m gpu -q 2>&1 | tail -30
python bench/har_step.py 2>&1 | tail -3 | tee gpurun_out/har_step.json
