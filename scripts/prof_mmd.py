"""One full-size MMD call per kernel type for ncu (argv: N, kernel type 0|1, passes)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
kt = int(sys.argv[2]) if len(sys.argv) > 2 else 0
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 0
g = torch.Generator(device="cuda").manual_seed(1265)
z = torch.randn(N, 128, device="cuda", generator=g) + 0.25; p = torch.randn(N, 128, device="cuda", generator=g)
for _ in range(2):
    F.mmd_sweep(z, p, kt, 2.0, 0, N, passes)
torch.cuda.synchronize()
