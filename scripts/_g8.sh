#!/bin/bash
# multi-GPU session (gpurun --gpus N): the driver's own command at every rank count up to N, the NCCL parity tests, summary
#   gpurun --gpus 8 --timeout 1500 -- 'bash scripts/_g8.sh r02e 2 4 8'
TAG=${1:-r02e}; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py -q -rA > gpurun_out/pytest_nccl_$TAG.log 2>&1
echo "nccl pytest rc=$? $(grep -E 'passed|failed|skipped' gpurun_out/pytest_nccl_$TAG.log | tail -1)"
for n in "$@"; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/scale_all_${TAG}_n$n.json 2> gpurun_out/scale_all_${TAG}_n$n.err
  echo "n=$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_all_${TAG}_n$n.json").read().strip().splitlines()[-1])
    rows=[(d["headline_workload"], d)] + list(d["workloads"].items())
    for k,v in rows:
        if "error" in v: print("  ", k, "ERROR", v["error"][:200]); continue
        e=v.get("e2e") or {}
        print("   %-12s %8.4f ms  value %14.1f  e2e %.3f ms  roofline %s %.3f  %s" % (k, v["ms_per_step"], v["value"], e.get("ms_per_step",0), v["roofline"].get("kernel"), v["roofline"].get("frac",0), v["scaling"]))
except Exception as e:
    print("   failed", e); print(open("gpurun_out/scale_all_${TAG}_n$n.err").read()[-2500:])
PY
done
