#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_native_train.py tests/test_engine.py -m gpu -q 2>&1 | tail -5
python bench.py --steps 5 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_resident.json | cut -c1-330
ATTACKFL_RESIDENT=0 python bench.py --steps 5 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_nonresident.json | cut -c1-330
