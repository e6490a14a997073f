#!/bin/bash
# r01j: full GPU parity suite, bench lines of every workload, reference arm
TAG=r01j
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; tail -3 gpurun_out/pytest_gpu_${TAG}.log
python bench.py > gpurun_out/bench_iwae_${TAG}.json 2> gpurun_out/bench_iwae_${TAG}.err; cut -c1-400 gpurun_out/bench_iwae_${TAG}.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2>&1; cut -c1-400 gpurun_out/bench_ref_${TAG}.json
for w in wae_mmd_rbf wae_mmd_imq vq swae betatc; do
  python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err
  tail -c 300 gpurun_out/bench_${w}_${TAG}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${w}_${TAG}.json").read().strip().splitlines()[-1])
print("${w}", d["ms_per_step"], d["value"], d["roofline"]["frac"], {k:round(v["ms"],4) for k,v in d["roofline"]["calls"].items()})
PY
done
for w in vq swae; do bash scripts/ncu_list.sh $w $TAG; python scripts/summarize_launches.py gpurun_out/launches_${w}_${TAG}.csv | tail -14; done
