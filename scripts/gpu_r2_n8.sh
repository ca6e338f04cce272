#!/usr/bin/env bash
# Round-2 8-GPU session: headline bench, configs 2 and 4 (ours; reference arm for config 2), NVLS kernel test.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29541 bench.py --gpus 8 --steps 10 --warmup 3 2>gpurun_out/bench_r2_n8.err | tail -1 > gpurun_out/bench_r2_n8.json
$TR --master-port 29542 bench/configs.py --config 2 --rounds 10 --warmup 3 2>gpurun_out/cfg2_n8.err | tail -1 > gpurun_out/cfg2_r2_n8.json
$TR --master-port 29543 bench/configs.py --config 2 --impl reference --rounds 2 --warmup 1 2>gpurun_out/cfg2_ref_n8.err | tail -1 > gpurun_out/cfg2_r2_ref_n8.json
$TR --master-port 29544 bench/configs.py --config 4 --rounds 3 --warmup 2 2>gpurun_out/cfg4_n8.err | tail -1 > gpurun_out/cfg4_r2_n8.json
timeout 400 python -m pytest "tests/test_distributed.py::test_fused_kernel_against_torch_and_flag_stress[8]" -x -q 2>&1 | tail -3 > gpurun_out/test_kernel_n8.log
for f in bench_r2_n8 cfg2_r2_n8 cfg2_r2_ref_n8 cfg4_r2_n8; do echo "== $f"; cut -c1-900 gpurun_out/$f.json; done
cat gpurun_out/test_kernel_n8.log
tail -2 gpurun_out/*.err | head -30
