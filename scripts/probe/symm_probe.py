"""torchrun --nproc-per-node 2 scripts/probe/symm_probe.py: can this box give us peer-mapped buffers
(torch.distributed._symmetric_memory) and can libvaekit kernels store into / load from them?"""
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
from vaekit import functional as F  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
n, D, S = 1024, 128, 16
t = symm.empty((S, world * n), dtype=torch.float32, device=dev)
hdl = symm.rendezvous(t, dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast", hdl.has_multicast_support, flush=True)
g = torch.Generator(device=dev).manual_seed(rank)
x = torch.randn(n, D, device=dev, generator=g)
proj = torch.randn(S, D, generator=torch.Generator().manual_seed(1)).to(dev)
t.zero_()
hdl.barrier(channel=0)
for peer in range(world):
    view = hdl.get_buffer(peer, (S, world * n), torch.float32)
    F.swd_project(x, proj, out=view[:, rank * n:(rank + 1) * n])       # tcgen05 GEMM epilogue storing into peer memory
hdl.barrier(channel=0)
torch.cuda.synchronize()
xs = [torch.empty_like(x) for _ in range(world)]
dist.all_gather(xs, x)
ref = torch.cat([xx @ proj.t() for xx in xs]).t()
print(rank, "max err of the peer-written projections", (t - ref).abs().max().item(), flush=True)
# loads from peer memory: gradient-GEMM style read of a peer's rows
peer = (rank + 1) % world
pv = hdl.get_buffer(peer, (S, world * n), torch.float32)
back = F.swd_project_back(pv[:, :n], proj)
print(rank, "peer-read GEMM err", (back - ref[:, :n].t() @ proj).abs().max().item(), flush=True)
hdl.barrier(channel=0)
dist.destroy_process_group()
