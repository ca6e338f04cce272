#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
python bench/kernels.py --json gpurun_out/kernels.json 2>&1 | tail -60 | tee gpurun_out/kernels.log
# ncu: gram kernel and fedavg kernel on mid-size shapes (plain run above exited 0 with the same code)
python bench/kernels.py --only gram > gpurun_out/gram_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gram_kernel -s 12 -c 1 -o gpurun_out/prof_gram -f python bench/kernels.py --only gram > gpurun_out/ncu_gram.log 2>&1
python bench/kernels.py --only fedavg > gpurun_out/fedavg_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fedavg_kernel -s 90 -c 1 -o gpurun_out/prof_fedavg -f python bench/kernels.py --only fedavg > gpurun_out/ncu_fedavg.log 2>&1
tail -2 gpurun_out/ncu_gram.log gpurun_out/ncu_fedavg.log
