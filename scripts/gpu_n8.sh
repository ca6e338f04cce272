#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2970$n bench.py --gpus $n --steps 5 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_n${n}_latest.json | cut -c1-260
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29718 bench.py --gpus 8 --steps 1 --warmup 3 --impl reference 2>&1 | tail -1 | tee gpurun_out/bench_ref_n8_latest.json | cut -c1-260
