#!/bin/bash
# A/B of the MMD gradient accumulation: default (chains cut by scheduler items, validated r01z) against
# -DVK_MMD_INNER_DRAIN=1 (chains cut inside an item; not validated yet).  Run on a GPU box:
#   gpurun --timeout 300 -- 'bash scripts/mmd_inner_ab.sh'
# Builds the variant into build_inner/ (nvcc is on the box), runs the MMD parity suite on it under a timeout (a barrier
# mistake would hang, not crash), then one bench line per library.
set -u
mkdir -p gpurun_out
P=$PWD/pytorch-vae_b200
VK_OBJ=$P/build_inner VK_OUT=$P/build_inner/libvaekit_inner.so VK_NVCC_EXTRA="-DVK_MMD_INNER_DRAIN=1" python $P/build.py | tail -1
VAEKIT_LIB=$P/build_inner/libvaekit_inner.so timeout 600 python -m pytest tests/test_gpu_mmd.py -q -rA > gpurun_out/pytest_mmd_inner.log 2>&1
echo "inner-drain pytest rc=$?" | tee -a gpurun_out/pytest_mmd_inner.log
grep -E "N=65536|N=4096|passed|failed" gpurun_out/pytest_mmd_inner.log | tail -8
for v in default inner; do
  lib=$P/vaekit/_lib/libvaekit.so; [ $v = inner ] && lib=$P/build_inner/libvaekit_inner.so
  VAEKIT_LIB=$lib timeout 200 python bench.py --workload wae_mmd_rbf --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_mmd_$v.json 2>> gpurun_out/mmd_inner_err.log
  echo "$v: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_mmd_$v.json)"
done
VAEKIT_LIB=$P/build_inner/libvaekit_inner.so timeout 200 python bench.py --workload wae_mmd_imq --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_mmd_imq_inner.json 2>> gpurun_out/mmd_inner_err.log
echo "imq inner: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_mmd_imq_inner.json)"
