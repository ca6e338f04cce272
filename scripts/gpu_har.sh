#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_har_native.py -m gpu -q 2>&1 | tail -15
