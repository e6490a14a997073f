"""The two SWAE GEMM shapes through vk_gemm_nt, for ncu captures of gemm3x_kernel."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
N = 1 << 20
z = torch.randn(N, 128, device="cuda"); th = torch.randn(200, 128, device="cuda")
g = torch.randn(200, N, device="cuda")
for _ in range(2):
    c = F.gemm_nt(z, th)                 # projection: row-major A
    d = F.gemm_nt(g.t(), th.t())         # gradient: k-major A (strided view), J = 128, K = 200
torch.cuda.synchronize()
print("ok", c.shape, d.shape)
