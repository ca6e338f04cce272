#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
: > gpurun_out/configs.log
run() { echo "== $*" | tee -a gpurun_out/configs.log; timeout 1500 python bench/configs.py "$@" 2>&1 | grep '^{' | tail -1 | tee -a gpurun_out/configs.log; }
run --config 2 --rounds 3 --warmup 3
run --config 4 --attack Min-Max --rounds 3 --warmup 3
run --config 4 --attack Min-Sum --rounds 2 --warmup 2
run --config 4 --attack Opt-Fang --rounds 2 --warmup 2
run --config 4 --attack Min-Max --clients 16 --rounds 3 --warmup 3
run --config 2 --rounds 1 --warmup 1 --impl reference
run --config 4 --attack Min-Max --clients 16 --rounds 1 --warmup 2 --impl reference
