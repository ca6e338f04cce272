#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_native_ops.py tests/test_engine.py -m gpu -q -k "select or defences" 2>&1 | tail -4
python bench/kernels.py --only select 2>&1 | tail -8 | tee gpurun_out/kernels_select.log
python bench/kernels.py --only gram 2>&1 | grep tcgen05 | tee gpurun_out/kernels_gram_tf32.log
# ncu of the tf32 gram kernel on the largest stock shape (R=256, D=203,776) after the same command exited 0
python bench/kernels.py --only gram > gpurun_out/gram_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gram_kernel -s 39 -c 1 -o gpurun_out/prof_gram_tf32 -f python bench/kernels.py --only gram > gpurun_out/ncu_gram_tf32.log 2>&1
tail -n 2 gpurun_out/ncu_gram_tf32.log
