#!/usr/bin/env bash
# Round-2 N-GPU session (N = 2 or 4): configs 2 and 4 with this repo's arm (config 2 also with the reference arm, one
# timed round), the slice-owned ScionFL / median defences, the sharded-storage hyper test and one fused-vs-NCCL run.
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29551 bench/configs.py --config 2 --rounds 10 --warmup 3 2>gpurun_out/cfg2_n$N.err | tail -1 > gpurun_out/cfg2_r2_n$N.json
$TR --master-port 29552 bench/configs.py --config 4 --rounds 3 --warmup 2 2>gpurun_out/cfg4_n$N.err | tail -1 > gpurun_out/cfg4_r2_n$N.json
if [ "${REF:-1}" = "1" ]; then
$TR --master-port 29553 bench/configs.py --config 2 --impl reference --rounds 1 --warmup 1 2>gpurun_out/cfg2_ref_n$N.err | tail -1 > gpurun_out/cfg2_r2_ref_n$N.json
fi
TESTS="tests/test_distributed.py::test_gpus_slice_owned_defences_match_gathered_path[scionfl-Random-$N] tests/test_distributed.py::test_gpus_slice_owned_defences_match_gathered_path[median-Random-$N]"
if [ "$N" = "2" ]; then TESTS="$TESTS tests/test_distributed.py::test_gpus_sharded_hyper_matches_gathered_path[2] tests/test_distributed.py::test_gpus_fused_matches_nccl_path[LIE-2]"; fi
if [ "$N" = "4" ]; then TESTS="$TESTS tests/test_distributed.py::test_gpus_fused_matches_nccl_path[Min-Max-4] tests/test_distributed.py::test_fused_kernel_against_torch_and_flag_stress[4]"; fi
timeout 900 python -m pytest $TESTS -x -q 2>&1 | tail -25 > gpurun_out/test_dist_n$N.log
for f in cfg2_r2_n$N cfg4_r2_n$N cfg2_r2_ref_n$N; do echo "== $f"; cut -c1-700 gpurun_out/$f.json 2>/dev/null; done
cat gpurun_out/test_dist_n$N.log
tail -2 gpurun_out/*_n$N.err | head -30
