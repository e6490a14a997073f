#!/bin/bash
TAG=r01n
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; tail -15 gpurun_out/pytest_gpu_${TAG}.log
for cfg in "1 peer" "2 all_to_all"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/prof_swd_dist.py $cfg 2>&1 | grep -v "^\*\|OMP_NUM\|^$\|warn\|Warn" | head -14 | tee gpurun_out/prof_swd_dist_n2_${cfg// /_}.log
done
