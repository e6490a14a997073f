"""torchrun --nproc-per-node N scripts/prof_swd_dist.py [waves]: GPU timeline summary of the all-to-all sharded SWD
(rank 0): per-kernel device time and the step's wall clock, to see how much of the exchange is exposed."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
from vaekit import dist as vd  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
waves = int(sys.argv[1]) if len(sys.argv) > 1 else None
exchange = sys.argv[2] if len(sys.argv) > 2 else "all_to_all"
N, D, S = 1 << 20, 128, 200
n = N // world
g = torch.Generator(device=dev).manual_seed(1265 + rank)
z = (1.5 * torch.randn(n, D, device=dev, generator=g) + 0.1).requires_grad_(True)
proj = torch.nn.functional.normalize(torch.randn(S, D, generator=torch.Generator().manual_seed(1)), dim=1).to(dev)
vd.shared_tensor_(proj)


def step():
    z.grad = None
    swd = vd.sharded_swd(z, proj, 2.0, 1e-10, exchange=exchange, waves=waves)
    swd.backward()
    return swd


for _ in range(3):
    step()
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    step()
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 5 * 1e3
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
if rank == 0:
    print("world %d waves %s exchange %s: %.3f ms/step (wall, 5 steps)" % (world, waves, exchange, ms))
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    agg = {}
    for e in ev:
        a = agg.setdefault(e.name[:70], [0, 0.0])
        a[0] += 1; a[1] += e.device_time
    for name, (cnt, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:18]:
        print("  %-70s n=%3d  %9.1f us/step" % (name, cnt // 3, us / 3))
    t_begin = min(e.time_range.start for e in ev); t_end = max(e.time_range.end for e in ev)
    print("  GPU span of 3 steps: %.3f ms/step; sum of kernel time %.3f ms/step" %
          ((t_end - t_begin) / 3e3, sum(v[1] for v in agg.values()) / 3e3))
dist.destroy_process_group()
