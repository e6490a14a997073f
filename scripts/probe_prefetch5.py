"""Which vaekit call serialises with a host->device copy on another stream?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
from vaekit import functional as F
dev = torch.device("cuda")
big = torch.randn(1 << 27).pin_memory(); dst = torch.empty(1 << 27, device=dev)
side = torch.cuda.Stream()
def copy():
    with torch.cuda.stream(side): dst.copy_(big, non_blocking=True)
def wall(fn, n=4):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
N, D, S = 1 << 20, 128, 200
g = torch.Generator(device=dev).manual_seed(1)
z = torch.randn(N, D, device=dev, generator=g); p = torch.randn(N, D, device=dev, generator=g)
th = torch.nn.functional.normalize(torch.randn(S, D, device=dev, generator=g), dim=1)
keys = torch.randn(S, N, device=dev, generator=g)
cases = {
    "gemm_nt x3": lambda: [F.gemm_nt(z, th) for _ in range(3)],
    "seg_sort keys": lambda: F.seg_sort(keys.clone()),
    "seg_sort keys+perm": lambda: F.seg_sort(keys.clone(), want_perm=True),
    "swd_sweep": lambda: F.swd_sweep(z, p, th, 2.0),
    "clone only x4": lambda: [keys.clone() for _ in range(4)],
}
c = wall(copy)
for name, fn in cases.items():
    a = wall(fn); b = wall(lambda: (copy(), fn()))
    print("%-20s alone %.2f ms | copy alone %.2f | both %.2f ms -> %s" % (name, a, c, b, "overlap" if b < 0.8 * (a + c) else "SERIAL"))
