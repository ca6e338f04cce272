#!/usr/bin/env bash
# Multi-GPU check: distributed tests + bench at N GPUs (both collective paths).
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m 2>&1 | head -12 > gpurun_out/topo.log
timeout 600 python -m pytest tests/test_distributed.py -m gpu -q -x 2>&1 | tail -15 | tee gpurun_out/pytest_dist.log
for coll in fused nccl; do
  echo "== bench N=$N collectives=$coll"
  ATTACKFL_COLLECTIVES=$coll timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 3 --warmup 3 2>&1 | tail -4 | tee gpurun_out/bench_n${N}_${coll}.log
done
