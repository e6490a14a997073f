#!/bin/bash
# MMD round-2 kernel: parity suite with the per-mode error table, chain-limit experiment for the IMQ, bench lines
set -u
mkdir -p gpurun_out
T=${1:-r02b}
timeout 900 python -m pytest tests/test_gpu_mmd.py -q -rA -x > gpurun_out/pytest_mmd_$T.log 2>&1
echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/pytest_mmd_$T.log | tail -2
grep -E "N=65536|auto:|unchecked|^FAILED|^ERROR|Error" gpurun_out/pytest_mmd_$T.log | head -60
for kt in rbf imq; do
  timeout 300 python bench.py --workload wae_mmd_$kt --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_mmd_${kt}_$T.json 2>> gpurun_out/bench_mmd_$T.err
  echo "$kt: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_mmd_${kt}_$T.json | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/bench_mmd_${kt}_$T.json | head -1)"
done
# IMQ accumulation chains: error / speed against the chain limit (0 = none)
for ch in 0 64 256; do
  VK_MMD_CHAIN=$ch timeout 300 python -m pytest "tests/test_gpu_mmd.py::test_mmd_full_size_vs_fp64_blocks[imq]" -q -rA 2>&1 | grep -E "passes=3|passes=0" | sed "s/^/chain=$ch /"
  VK_MMD_CHAIN=$ch timeout 200 python bench.py --workload wae_mmd_imq --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | grep -o '"ms_per_step": [0-9.]*' | head -1 | sed "s/^/chain=$ch /"
done
tail -3 gpurun_out/bench_mmd_$T.err
