"""Does a pinned host->device copy overlap an HBM-bound kernel stream (plain torch adds) / the IWAE streaming kernels?"""
import time, torch
dev = torch.device("cuda")
big = torch.randn(1 << 27).pin_memory(); dst = torch.empty(1 << 27, device=dev)
a = torch.randn(1 << 28, device=dev); b = torch.empty_like(a)
side = torch.cuda.Stream()
def copy():
    with torch.cuda.stream(side): dst.copy_(big, non_blocking=True)
def adds():
    for _ in range(30): torch.add(a, 1.0, out=b)
def wall(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("copy alone %.2f ms | 30 adds of 1 GB alone %.2f ms | both %.2f ms" % (wall(copy), wall(adds), wall(lambda: (copy(), adds()))))
def small():
    for _ in range(3000): torch.add(a[:1 << 16], 1.0, out=b[:1 << 16])
print("3000 tiny adds alone %.2f ms | + copy %.2f ms" % (wall(small), wall(lambda: (copy(), small()))))
