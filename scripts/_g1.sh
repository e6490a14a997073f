#!/bin/bash
# r01o: ncu --set full captures of the dominant kernel of every op workload (one launch each), + the new VQ backward
TAG=r01o
mkdir -p gpurun_out
cap() { # workload kernel-regex name skip
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$2" -s $4 -c 1 -f -o gpurun_out/ncu_$3_${TAG} \
    python bench.py --workload $1 --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_$3.log 2>&1; tail -1 gpurun_out/ncu_$3.log
}
cap vq vq_fwd_kernel vq_fwd 2
cap vq vq_bwd_kernel vq_bwd 2
cap swae "scatter_kernel<1, true" swae_scatter_kv 1
cap swae "scatter_kernel<1, false" swae_scatter_k 1
cap swae "scatter_kernel<3, true" swae_scatter_fused 1
cap wae_mmd_rbf mmd_main_kernel mmd_main 1
cap betatc bwd_rows_kernel tc_bwd_rows 1
ls -la gpurun_out/*.ncu-rep
