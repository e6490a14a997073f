import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
B, D = 4096, 128
mu = torch.randn(B, D, device="cuda"); lv = 0.3 * torch.randn(B, D, device="cuda")
z = (mu + torch.randn(B, D, device="cuda") * torch.exp(0.5 * lv))
zg, mg, lg = (t.clone().requires_grad_(True) for t in (z, mu, lv))
for _ in range(2):
    zg.grad = None; mg.grad = None; lg.grad = None
    F.betatc_terms(zg, mg, lg, 0.00025, 1.0, 6.0, 0.37)[0].backward()
torch.cuda.synchronize()
