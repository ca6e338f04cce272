#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_har_native.py -m gpu -q 2>&1 | tail -3
timeout 300 python bench/har_step.py 2>&1 | tail -1 | tee gpurun_out/har_step.json
timeout 600 python bench/configs.py --config 3 --clients 128 --rounds 2 --warmup 2 2>&1 | tail -1 | tee gpurun_out/cfg3_128.json | cut -c1-200
grep -o '"phases_ms.*' gpurun_out/cfg3_128.json | cut -c1-300
