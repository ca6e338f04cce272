#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_har_native.py -m gpu -q -k "tensor_core or grad_step" 2>&1 | tail -5
timeout 300 python bench/har_kernels.py 2>&1 | tail -1 | tee gpurun_out/har_kernels.json | cut -c1-330
