#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
: > gpurun_out/configs_extra.jsonl
for args in "--config 6 --rounds 2 --warmup 1" "--config 6 --rounds 1 --warmup 1 --backend torch"; do
  echo "== $args" >> gpurun_out/configs_extra.jsonl
  timeout 900 python bench/configs.py $args 2>&1 | tail -1 >> gpurun_out/configs_extra.jsonl
done
cut -c1-600 gpurun_out/configs_extra.jsonl
