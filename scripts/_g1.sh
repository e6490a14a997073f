#!/bin/bash
# r01q: full GPU parity suite, smoke, bench lines of every workload (value pass uninstrumented), reference arm
TAG=r01q
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; tail -3 gpurun_out/pytest_gpu_${TAG}.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_iwae_${TAG}.json 2> gpurun_out/bench_iwae_${TAG}.err; cut -c1-300 gpurun_out/bench_iwae_${TAG}.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2>&1; cut -c1-200 gpurun_out/bench_ref_${TAG}.json
for w in miwae vanilla; do python bench.py --workload $w --steps 20 > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err; cut -c1-200 gpurun_out/bench_${w}_${TAG}.json; done
for w in wae_mmd_rbf wae_mmd_imq vq swae betatc; do
  python bench.py --workload $w --steps 20 --warmup 3 > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err
  tail -c 300 gpurun_out/bench_${w}_${TAG}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${w}_${TAG}.json").read().strip().splitlines()[-1])
print("${w}", d["ms_per_step"], d["value"], d["roofline"]["frac"], {k:round(v["ms"],4) for k,v in d["roofline"]["calls"].items()}, "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
PY
done
