#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_iwae_r01u_n2.json 2> gpurun_out/bench_iwae_r01u_n2.err
tail -c 400 gpurun_out/bench_iwae_r01u_n2.err; cut -c1-700 gpurun_out/bench_iwae_r01u_n2.json
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29503 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | cut -c1-300
