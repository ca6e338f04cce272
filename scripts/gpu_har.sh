#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_har_native.py -m gpu -q -k "tensor_core or grad_step" 2>&1 | tail -3
timeout 300 python bench/har_kernels.py 2>&1 | tail -1 | tee gpurun_out/har_kernels.json | cut -c1-330
python scripts/prof_attn.py && ncu --set full --clock-control none --import-source on -k regex:"attn_fwd_tc_kernel|attn_bwd_dq_tc_kernel" -s 2 -c 2 -o gpurun_out/prof_attn_fq -f python scripts/prof_attn.py > gpurun_out/ncu_attn_fq.log 2>&1; tail -1 gpurun_out/ncu_attn_fq.log
