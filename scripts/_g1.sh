timeout 600 python -m pytest tests/test_gpu_swd.py -x -q > gpurun_out/pytest_swd.log 2>&1; tail -3 gpurun_out/pytest_swd.log
timeout 300 python scripts/time_ops.py swd 2>&1 | tail -8
timeout 300 python bench.py --workload swae --steps 5 --warmup 3 --no-cpu > gpurun_out/b_swae.json 2>gpurun_out/b_swae.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_swae_r01j.csv python bench.py --workload swae --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1; python scripts/summarize_launches.py gpurun_out/launches_swae_r01j.csv 2>&1 | head -30
