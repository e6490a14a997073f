#!/bin/bash
# usage: ncu_list.sh <workload> <tag>  -- launch list (per-kernel device time) of one bench workload
w=$1; TAG=$2
CMD="python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_${w}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_${w}_${TAG}.csv $CMD > gpurun_out/ncu_list_${w}.log 2>&1
