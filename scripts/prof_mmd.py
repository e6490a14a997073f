import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
z = torch.randn(N, 128, device="cuda") + 0.25; p = torch.randn(N, 128, device="cuda")
for _ in range(2):
    F.mmd_sweep(z, p, 0, 2.0, 0, N, 3)
torch.cuda.synchronize()
