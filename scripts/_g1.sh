#!/bin/bash
# scratch: r01z -- MMD accumulation chains cut at 32 tiles: parity at full size + step time; SWD full-size stages
mkdir -p gpurun_out
timeout 45 python -m pytest tests/test_gpu_mmd.py tests/test_gpu_swd.py::test_swd_full_size_stages -q -rA > gpurun_out/pytest_fullsize_r01z.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_fullsize_r01z.log
grep -E "N=65536|N=4096|full-size|passed|failed|^E  |rc=" gpurun_out/pytest_fullsize_r01z.log | head -24
timeout 25 python bench.py --workload wae_mmd_rbf --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_wae_mmd_rbf_r01z.json 2> gpurun_out/r01z_err.log
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_wae_mmd_rbf_r01z.json
