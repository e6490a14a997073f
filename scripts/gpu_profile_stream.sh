#!/bin/bash
# Runs on the B200 box (via gpurun): ncu launch list + one full capture of the IW-bound kernels.
# Usage: bash scripts/gpu_profile_stream.sh <round-tag>
TAG=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_list_${TAG}.log 2>&1
$CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:iw_bound -s 2 -c 2 \
    -o gpurun_out/prof_iw_${TAG} $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
ls -la gpurun_out
