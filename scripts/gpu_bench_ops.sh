#!/bin/bash
# Runs on the B200 box: bench lines of every op workload (+ IWAE default) and the ncu launch lists.
TAG=${1:-r01c}
mkdir -p gpurun_out
for w in wae_mmd_rbf wae_mmd_imq vq swae betatc; do
  python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err
  tail -c 600 gpurun_out/bench_${w}_${TAG}.err
done
python bench.py > gpurun_out/bench_iwae_${TAG}.json 2> gpurun_out/bench_iwae_${TAG}.err
for w in wae_mmd_rbf vq swae betatc; do
  CMD="python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu"
  $CMD > gpurun_out/plain_${w}.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_${w}_${TAG}.csv $CMD > gpurun_out/ncu_list_${w}.log 2>&1
done
cat gpurun_out/bench_*_${TAG}.json | cut -c1-1500
