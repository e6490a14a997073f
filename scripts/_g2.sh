#!/bin/bash
TAG=r01p
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_betatc.py tests/test_gpu_dist_nccl.py tests/test_gpu_swd.py -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; tail -25 gpurun_out/pytest_gpu_${TAG}.log
