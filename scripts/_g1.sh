#!/bin/bash
# one-GPU session: GPU test suite, then the default bench line
set -u
mkdir -p gpurun_out
T=${1:-r02a}
timeout 1200 python -m pytest tests -m gpu -q -rA > gpurun_out/pytest_gpu_$T.log 2>&1
echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/pytest_gpu_$T.log | tail -3; grep -E "^FAILED|^ERROR" gpurun_out/pytest_gpu_$T.log | head -20
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err
echo "bench rc=$?"; tail -5 gpurun_out/bench_$T.err; head -c 1500 gpurun_out/bench_$T.json
