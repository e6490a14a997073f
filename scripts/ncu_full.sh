#!/bin/bash
# usage: ncu_full.sh <prof script> <kernel regex> <tag>  -- one --set full capture (with source) of the named kernel
S=$1; K=$2; TAG=$3
python $S > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$K" -s ${SKIP:-1} -c 1 -f -o gpurun_out/${TAG} python $S > gpurun_out/ncu_${TAG}.log 2>&1
