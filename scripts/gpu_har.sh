#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_har_native.py -m gpu -q -x -k "tensor_core" 2>&1 | tail -30
