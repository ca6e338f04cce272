#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29741 bench/configs.py --config 2 --rounds 3 --warmup 3 2>&1 | tail -1 | tee gpurun_out/cfg2_n8.json | cut -c1-700
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29742 bench/configs.py --config 4 --attack Min-Max --rounds 3 --warmup 3 2>&1 | tail -1 | tee gpurun_out/cfg4_n8.json | cut -c1-200
grep -o '"phases_ms.*' gpurun_out/cfg4_n8.json | cut -c1-300
