#!/bin/bash
# MMD profile session: launch list of the bench workload, one --set full capture of the main kernel per kernel type
set -u
mkdir -p gpurun_out
T=${1:-r02d}
timeout 300 python bench.py --workload vanilla_b64 --steps 10 --warmup 3 > gpurun_out/bench_vanilla_b64_$T.json 2> gpurun_out/bench_vanilla_b64_$T.err
echo "vanilla_b64 rc=$?"; tail -2 gpurun_out/bench_vanilla_b64_$T.err | cut -c1-300; head -c 900 gpurun_out/bench_vanilla_b64_$T.json; echo
for w in wae_mmd_rbf wae_mmd_imq; do
  timeout 600 bash scripts/ncu_list.sh $w $T
  python scripts/summarize_launches.py gpurun_out/launches_${w}_$T.csv 12
done
SKIP=1 timeout 900 bash scripts/ncu_full.sh "scripts/prof_mmd.py 65536 0 0" "mmd_main_kernel" ncu_mmd_rbf_$T
python scripts/ncu_metrics.py gpurun_out/ncu_mmd_rbf_$T.ncu-rep > gpurun_out/ncu_mmd_rbf_${T}_summary.txt 2>&1; head -60 gpurun_out/ncu_mmd_rbf_${T}_summary.txt
