#!/bin/bash
# r01v: ncu --set full of the SWD sort kernels (one key-only scatter, one key+index scatter, the fused last pass, one histogram) and the tcgen05 GEMM
mkdir -p gpurun_out
CMD="python bench.py --workload swae --steps 1 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/plain_swae.log 2>&1
# launch order per step: gemm x2, then prior sort (hist,scan,scatter)x4, then z sort x4 (last fused), gemm
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:scatter_kernel|hist_kernel|gemm3x_kernel" -s 57 -c 19 -f -o gpurun_out/ncu_swae_r01v $CMD > gpurun_out/ncu_swae.log 2>&1; tail -2 gpurun_out/ncu_swae.log
ls -la gpurun_out/ncu_swae_r01v.ncu-rep
