"""CUDA-event timing of vk_mmd_fwd per precision mode at the benchmark size (argv: N)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
g = torch.Generator(device="cuda").manual_seed(1265)
z = torch.randn(N, 128, device="cuda", generator=g) + 0.25; p = torch.randn(N, 128, device="cuda", generator=g)
for kt, modes in ((0, (1, 2, 3, 0)), (1, (3, 22, 31, 32, 0))):
    for passes in modes:
        for _ in range(3):
            F.mmd_sweep(z, p, kt, 2.0, 0, N, passes)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            F.mmd_sweep(z, p, kt, 2.0, 0, N, passes)
        e1.record(); torch.cuda.synchronize()
        print("kernel_type %d passes %d: %.3f ms per vk_mmd_fwd" % (kt, passes, e0.elapsed_time(e1) / 10), flush=True)
