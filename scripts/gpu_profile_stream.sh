#!/bin/bash
# Runs on the B200 box (via gpurun): GPU tests, bench, ncu launch list + one full capture of the IW-bound kernels.
# Usage: bash scripts/gpu_profile_stream.sh <round-tag>
TAG=${1:-r01}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_${TAG}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${TAG}.log
python bench.py > gpurun_out/bench_iwae_${TAG}.json 2> gpurun_out/bench_iwae_${TAG}.err
python bench.py --workload miwae --steps 20 > gpurun_out/bench_miwae_${TAG}.json 2> gpurun_out/bench_miwae_${TAG}.err
python bench.py --workload vanilla --steps 20 > gpurun_out/bench_vanilla_${TAG}.json 2> gpurun_out/bench_vanilla_${TAG}.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err
CMD="python bench.py --steps 3 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_list_${TAG}.log 2>&1
$CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:iw_bound -s 2 -c 2 \
    -o gpurun_out/prof_iw_${TAG} $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
ls -la gpurun_out
tail -3 gpurun_out/pytest_${TAG}.log; cat gpurun_out/bench_*_${TAG}.json
