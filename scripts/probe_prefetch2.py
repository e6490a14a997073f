"""Why does the e2e leg of bench.py not overlap the next batch's copy with the step?  Times the SWAE and MMD steps with the
prefetcher, with host-side timestamps."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
import bench
from vaekit.io import DevicePrefetcher
dev = torch.device("cuda")
for name in sys.argv[1:] or ["swae"]:
    wl = bench.WORKLOADS[name]()
    model = wl.model(dev)
    if isinstance(model, torch.nn.Module): model = model.to(dev)
    t = wl.make(dev, 0, 1)
    host = {k: v.detach().cpu().pin_memory() for k, v in t.items()}
    for _ in range(3): wl.step(model, t, 1)
    torch.cuda.synchronize()
    def wall(fn, n=5):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n): fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e3
    print(name, "step only            %.2f ms" % wall(lambda: [v.item() for v in wl.step(model, t, 1).values()]))
    print(name, "copy then step       %.2f ms" % wall(lambda: [v.item() for v in wl.step(model, {k: v.to(dev, non_blocking=True) for k, v in host.items()}, 1).values()]))
    pf = DevicePrefetcher(dev); pf.submit(host)
    def pstep():
        td = pf.get(); pf.submit(host)
        h0 = time.perf_counter()
        out = wl.step(model, td, 1)
        h1 = time.perf_counter()
        r = [v.item() for v in out.values()]
        pstep.host_ms = (h1 - h0) * 1e3
        return r
    print(name, "prefetch + step      %.2f ms  (host time inside step() %.2f ms)" % (wall(pstep), pstep.host_ms))
    side = torch.cuda.Stream()
    dst = {k: torch.empty_like(v, device=dev) for k, v in host.items()}
    def raw():
        with torch.cuda.stream(side):
            for k, v in host.items(): dst[k].copy_(v, non_blocking=True)
        return [v.item() for v in wl.step(model, t, 1).values()]
    print(name, "unrelated copy + step %.2f ms" % wall(raw))
