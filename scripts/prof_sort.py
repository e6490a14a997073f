"""Two segmented sorts (key-only, then key + permutation) for ncu captures of the radix-sort kernels."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
S, N = int(os.environ.get("S", 50)), 1 << 20
keys = torch.randn(S, N, device="cuda")
for _ in range(2):
    F.seg_sort(keys)
    F.seg_sort(keys, want_perm=True)
torch.cuda.synchronize()
print("ok")
