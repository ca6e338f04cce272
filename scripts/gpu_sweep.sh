#!/usr/bin/env bash
# Aggregation sweep (config 5) + headline bench at N GPUs.
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547 bench/agg_sweep.py --rows 1 2>&1 | grep '^{' | tee gpurun_out/agg_sweep_n${N}.jsonl
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29548 bench/agg_sweep.py --rows 8 --sizes-mb 1,16,128 2>&1 | grep '^{' | tee gpurun_out/agg_sweep_n${N}_rows8.jsonl
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29549 bench.py --gpus $N --steps 3 --warmup 3 2>&1 | grep '^{' | tee gpurun_out/bench_n${N}.json
