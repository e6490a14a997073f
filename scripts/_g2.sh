#!/bin/bash
TAG=r01s
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_betatc.py tests/test_gpu_dist_nccl.py tests/test_gpu_next_models.py -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; tail -12 gpurun_out/pytest_gpu_${TAG}.log
VAEKIT_TC_NO_FWD_SPLIT=1 timeout 300 python -m pytest tests/test_gpu_betatc.py -x -q -k "shards" 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --workload betatc --gpus 2 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/scale_betatc_${TAG}_n2.json 2> gpurun_out/scale_betatc_${TAG}_n2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/scale_betatc_${TAG}_n2.json").read().strip().splitlines()[-1])
print("betatc n=2", d["ms_per_step"], d["value"], {k:round(v["ms"],4) for k,v in d["roofline"]["calls"].items()})
PY
