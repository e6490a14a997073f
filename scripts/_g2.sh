#!/bin/bash
# MMD session: parity suite with the per-mode error table, bench lines, launch list, one --set full capture
set -u
mkdir -p gpurun_out
T=${1:-r02e}
timeout 900 python -m pytest tests/test_gpu_mmd.py -q -rA -x > gpurun_out/pytest_mmd_$T.log 2>&1
echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/pytest_mmd_$T.log | tail -2
grep -E "N=65536|^FAILED|^ERROR|Error" gpurun_out/pytest_mmd_$T.log | head -20
for kt in rbf imq; do
  timeout 300 python bench.py --workload wae_mmd_$kt --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_mmd_${kt}_$T.json 2>> gpurun_out/bench_mmd_$T.err
  echo "$kt: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_mmd_${kt}_$T.json | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/bench_mmd_${kt}_$T.json | head -1)"
done
timeout 600 bash scripts/ncu_list.sh wae_mmd_rbf $T
python scripts/summarize_launches.py gpurun_out/launches_wae_mmd_rbf_$T.csv 8
SKIP=1 timeout 900 bash scripts/ncu_full.sh "scripts/prof_mmd.py 65536 0 0" "mmd_main_kernel<2" ncu_mmd_rbf_$T
python scripts/ncu_metrics.py gpurun_out/ncu_mmd_rbf_$T.ncu-rep > gpurun_out/ncu_mmd_rbf_${T}_summary.txt 2>&1; head -40 gpurun_out/ncu_mmd_rbf_${T}_summary.txt
tail -3 gpurun_out/bench_mmd_$T.err
