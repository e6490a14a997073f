#!/bin/bash
mkdir -p gpurun_out
timeout 28 python -m pytest tests/test_gpu_swd.py::test_swd_full_size_stages -q -rA > gpurun_out/pytest_swd_fullsize_r01z2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_swd_fullsize_r01z2.log
grep -E "full-size|passed|failed|^E  |rc=" gpurun_out/pytest_swd_fullsize_r01z2.log | head -20
