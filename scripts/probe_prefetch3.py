"""Which part of a step serialises with a host->device copy on another stream?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
import bench
dev = torch.device("cuda")
wl = bench.WORKLOADS["swae"]()
model = wl.model(dev).to(dev)
t = wl.make(dev, 0, 1)
host = {k: v.detach().cpu().pin_memory() for k, v in t.items()}
dst = {k: torch.empty_like(v, device=dev) for k, v in host.items()}
side = torch.cuda.Stream()
for _ in range(3): wl.step(model, t, 1)
torch.cuda.synchronize()
def wall(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
def copy():
    with torch.cuda.stream(side):
        for k, v in host.items(): dst[k].copy_(v, non_blocking=True)
def fwd_only():
    z = t["z"].detach()
    n = z.shape[0]
    with torch.no_grad():
        return model.compute_swd(z, model.p, model.reg_weight / (n * (n - 1)))
def full_noitem():
    return wl.step(model, t, 1)
def sync_each(fn):
    def g():
        r = fn(); torch.cuda.current_stream().synchronize(); return r
    return g
print("copy alone                  %.2f ms" % wall(copy))
for name, fn in (("forward only", fwd_only), ("full step, no item", full_noitem)):
    print("%-27s %.2f ms | + copy %.2f ms | + copy, stream sync each step %.2f ms" % (name, wall(fn), wall(lambda: (copy(), fn())), wall(sync_each(lambda: (copy(), fn())))))
print("full step + item + copy     %.2f ms" % wall(lambda: (copy(), [v.item() for v in wl.step(model, t, 1).values()])))
