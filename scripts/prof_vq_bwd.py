"""Forward + backward of the VectorQuantizer at the benchmark size, three times (for ncu: -k regex:vq_bwd_kernel)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
g = torch.Generator(device="cuda").manual_seed(1265)
lat = torch.nn.functional.leaky_relu(torch.randn(1024, 64, 16, 16, device="cuda", generator=g)).requires_grad_(True)
cb = torch.randn(512, 64, device="cuda", generator=g).requires_grad_(True)
gq = 1e-3 * torch.randn(1024, 64, 16, 16, device="cuda", generator=g)
for _ in range(3):
    lat.grad = None; cb.grad = None
    q, l, idx = F.vector_quantize(lat, cb, 0.25)
    torch.autograd.backward([q, l], [gq, None])
torch.cuda.synchronize()
