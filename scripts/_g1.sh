#!/bin/bash
# scratch: r01x -- e2e of the ops workloads after the host-copy fix (no CPU baseline leg, 10 steps)
mkdir -p gpurun_out
for w in swae vq betatc wae_mmd_rbf; do
  timeout 60 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_${w}_r01x.json 2>> gpurun_out/r01x_err.log
  python - gpurun_out/bench_${w}_r01x.json <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); print(sys.argv[1], d['ms_per_step'], d['e2e'], d['roofline'].get('traffic'))
PY
done
