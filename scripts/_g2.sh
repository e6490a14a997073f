#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/probe/symm_probe.py 2>&1 | grep -v "^\*\|OMP_NUM\|^$" | tail -30 | tee gpurun_out/symm_probe.log
nvidia-smi topo -m | head -12
