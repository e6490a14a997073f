#!/bin/bash
TAG=${1:-r01t}
mkdir -p gpurun_out
run() { # workload n
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $((29500 + $2)) \
    bench.py --workload $1 --gpus $2 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/scale_$1_${TAG}_n$2.json 2> gpurun_out/scale_$1_${TAG}_n$2.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_$1_${TAG}_n$2.json").read().strip().splitlines()[-1])
    print("$1 n=$2", d["ms_per_step"], d["value"], d["roofline"]["frac"], {k:round(v["ms"],4) for k,v in d["roofline"]["calls"].items()})
except Exception as e:
    print("$1 n=$2 failed", e); print(open("gpurun_out/scale_$1_${TAG}_n$2.err").read()[-1500:])
PY
}
run betatc 8
run betatc 4
