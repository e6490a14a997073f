#!/bin/bash
# scratch: r01w regression -- full GPU suite, smoke, SWAE bench line, launch list
mkdir -p gpurun_out
timeout 150 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_r01w.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r01w.log
tail -3 gpurun_out/pytest_gpu_r01w.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_r01w.log 2>&1; tail -1 gpurun_out/smoke_r01w.log
timeout 90 python bench.py --workload swae > gpurun_out/bench_swae_r01w.json 2> gpurun_out/bench_swae_r01w.err; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_swae_r01w.json
timeout 70 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_list_swae_r01w.csv python bench.py --workload swae --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_list_swae_r01w.log 2>&1; echo "ncu rc=$?"
