#!/bin/bash
# r01u: full GPU parity suite + smoke + short bench lines after the host fast paths / Gumbel kernels
TAG=r01u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; tail -15 gpurun_out/pytest_gpu_${TAG}.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_iwae_${TAG}.json 2> gpurun_out/bench_iwae_${TAG}.err; cut -c1-200 gpurun_out/bench_iwae_${TAG}.json; tail -c 300 gpurun_out/bench_iwae_${TAG}.err
for w in vq betatc; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err
  tail -c 300 gpurun_out/bench_${w}_${TAG}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${w}_${TAG}.json").read().strip().splitlines()[-1])
print("${w}", d["ms_per_step"], d["value"], d["roofline"]["frac"], {k:round(v["ms"],4) for k,v in d["roofline"]["calls"].items()})
PY
done
python scripts/host_prof_vq.py vq 2>&1 | head -3
