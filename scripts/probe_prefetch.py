"""Does a pinned H2D copy on a side stream overlap a running step?  (copy alone, step alone, both)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
dev = torch.device("cuda")
N = 65536
z = torch.randn(N, 128, device=dev) + 0.25; p = torch.randn(N, 128, device=dev)
big = torch.randn(1 << 27).pin_memory()            # 512 MB
dst = torch.empty(1 << 27, device=dev)
side = torch.cuda.Stream()
def step():
    for _ in range(2): F.mmd_sweep(z, p, 0, 2.0, 0, N, 2)
def copy():
    with torch.cuda.stream(side): dst.copy_(big, non_blocking=True)
def wall(fn, n=5):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3
print("copy alone  %.2f ms" % wall(copy))
print("step alone  %.2f ms" % wall(step))
print("both        %.2f ms" % wall(lambda: (copy(), step())))
print("both (step first) %.2f ms" % wall(lambda: (step(), copy())))
