#!/bin/bash
# usage: gpu_scale.sh <tag> <N...>   -- sharded-op scaling bench (MMD rbf, SWAE) at the given GPU counts
TAG=$1; shift
for w in wae_mmd_rbf swae; do
  for n in "$@"; do
    if [ "$n" = 1 ]; then
      timeout 600 python bench.py --workload $w --gpus 1 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_${w}_${TAG}_n$n.json 2> gpurun_out/scale_${w}_${TAG}_n$n.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
        bench.py --workload $w --gpus $n --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_${w}_${TAG}_n$n.json 2> gpurun_out/scale_${w}_${TAG}_n$n.err
    fi
    tail -1 gpurun_out/scale_${w}_${TAG}_n$n.json
  done
done
