#!/bin/bash
TAG=r01l
mkdir -p gpurun_out
bash scripts/ncu_list.sh vq $TAG; python scripts/summarize_launches.py gpurun_out/launches_vq_${TAG}.csv | tail -12
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:vq_bwd_kernel" -s 2 -c 1 -f -o gpurun_out/vq_bwd_${TAG} python bench.py --workload vq --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_vq_bwd.log 2>&1; tail -2 gpurun_out/ncu_vq_bwd.log
