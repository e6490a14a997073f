#!/bin/bash
# MMD verification + profile session: parity suite, per-mode timing, launch list, one --set full capture of the (2,1) kernel
set -u
mkdir -p gpurun_out
T=${1:-r02g}
timeout 900 python -m pytest tests/test_gpu_mmd.py -q -rA -x > gpurun_out/pytest_mmd_$T.log 2>&1
echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/pytest_mmd_$T.log | tail -2
grep -E "N=65536|^FAILED|^ERROR|Error" gpurun_out/pytest_mmd_$T.log | head -20
python scripts/time_mmd.py 65536 2>&1 | tee gpurun_out/time_mmd_$T.log
timeout 600 bash scripts/ncu_list.sh wae_mmd_rbf $T
python scripts/summarize_launches.py gpurun_out/launches_wae_mmd_rbf_$T.csv 10 | tee gpurun_out/launches_wae_mmd_rbf_${T}_summary.txt
SKIP=2 timeout 900 bash scripts/ncu_full.sh "scripts/prof_mmd.py 65536 0 0" "mmd_main_kernel" ncu_mmd_rbf_$T
python scripts/ncu_metrics.py gpurun_out/ncu_mmd_rbf_$T.ncu-rep > gpurun_out/ncu_mmd_rbf_${T}_summary.txt 2>&1; head -32 gpurun_out/ncu_mmd_rbf_${T}_summary.txt
