#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 3 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_n2_latest.json | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 1 --warmup 3 --impl reference 2>&1 | tail -1 | tee gpurun_out/bench_ref_n2_latest.json | cut -c1-300
timeout 600 python -m pytest tests/test_distributed.py -m gpu -q 2>&1 | tail -3
