#!/usr/bin/env bash
# ncu --set full of the kernels added or changed late in round 2 (raw csv exported on the box: the .ncu-rep is large).
set -u
mkdir -p gpurun_out
timeout 300 python scripts/prof_kernels_r2.py > gpurun_out/prof_kernels_r2b.log 2>&1 || { tail -5 gpurun_out/prof_kernels_r2b.log; exit 1; }
tail -1 gpurun_out/prof_kernels_r2b.log
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'roc_auc_kernel|round_indices_kernel|tc_train_kernel|hyper_round_kernel|scion_scores_kernel|scion_quantize_kernel' \
  -o /tmp/r2b -f python scripts/prof_kernels_r2.py > gpurun_out/ncu_kernels_r2b.log 2>&1
tail -2 gpurun_out/ncu_kernels_r2b.log
ncu -i /tmp/r2b.ncu-rep --page raw --csv > gpurun_out/kernels_r2b_raw.csv 2>/dev/null
wc -c gpurun_out/kernels_r2b_raw.csv
