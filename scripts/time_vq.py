"""CUDA-event timing of vk_vq_fwd / vk_vq_bwd at BASELINE configs[3] (1024 x 64 x 16 x 16 latents, 512 x 64 codebook)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import _lib as L
from vaekit import functional as F
g = torch.Generator(device="cuda").manual_seed(1265)
lat = torch.nn.functional.leaky_relu(torch.randn(1024, 64, 16, 16, device="cuda", generator=g)).requires_grad_(True)
cb = torch.randn(512, 64, device="cuda", generator=g).requires_grad_(True)
gq = 1e-3 * torch.randn(1024, 64, 16, 16, device="cuda", generator=g)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def step():
    lat.grad = cb.grad = None
    q, l, _ = F.vector_quantize(lat, cb, 0.25)
    torch.autograd.backward([q, l], [gq, None])


for _ in range(3):
    step()
for cold in (False, True):
    L.EVENT_LOG = []
    for _ in range(10):
        if cold:
            flush.zero_()                 # 256 MB > L2: the next call starts from HBM
        step()
    torch.cuda.synchronize()
    acc = {}
    for name, a, b in L.EVENT_LOG:
        acc.setdefault(name, []).append(a.elapsed_time(b))
    L.EVENT_LOG = None
    print("L2 flushed between steps" if cold else "back to back", {k: "%.1f us" % (1e3 * sorted(v)[len(v) // 2]) for k, v in acc.items()}, flush=True)
