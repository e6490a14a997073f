"""Quick device timings of single ops (not the bench contract; used while tuning)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F

def timeit(fn, warm=2, it=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

dev = torch.device("cuda")
which = sys.argv[1:] or ["mmd", "vq"]
if "mmd" in which:
    for N in (8192, 65536):
        z = torch.randn(N, 128, device=dev) + 0.25; p = torch.randn(N, 128, device=dev)
        for kt in (0, 1):
            for passes in (3, 1):
                ms = timeit(lambda: F.mmd_sweep(z, p, kt, 2.0, 0, N, passes))
                fl = 10.0 * N * N * 128
                print("mmd N=%d kernel=%d passes=%d: %.3f ms  %.2f Msamples/s  algorithmic %.1f TFLOP/s" % (N, kt, passes, ms, N / ms / 1e3, fl / ms / 1e9), flush=True)
if "vq" in which:
    lat = torch.nn.functional.leaky_relu(torch.randn(1024, 64, 16, 16, device=dev)); cb = torch.randn(512, 64, device=dev)
    ms = timeit(lambda: F.vector_quantize(lat, cb, 0.25), it=20)
    print("vq fwd B=1024: %.4f ms  %.2f M images/s, %.1f GB/s on 135.4 MB" % (ms, 1024 / ms / 1e3, 135.4e6 / ms / 1e6))
    latg = lat.clone().requires_grad_(True); cbg = cb.clone().requires_grad_(True)
    gq = torch.randn_like(lat) * 1e-3
    def step():
        latg.grad = None; cbg.grad = None
        q, l, _ = F.vector_quantize(latg, cbg, 0.25)
        torch.autograd.backward([q, l], [gq, None])
    ms = timeit(step, it=20)
    print("vq fwd+bwd B=1024: %.4f ms  %.2f M images/s, %.1f GB/s on 337.9 MB" % (ms, 1024 / ms / 1e3, 337.9e6 / ms / 1e6))
if "tc" in which:
    B, D = 4096, 128
    mu = torch.randn(B, D, device=dev); lv = 0.3 * torch.randn(B, D, device=dev)
    z = mu + torch.randn(B, D, device=dev) * torch.exp(0.5 * lv)
    ms = timeit(lambda: F.betatc_terms(z, mu, lv, 0.00025, 1.0, 6.0, 0.37), warm=1, it=3)
    print("betatc fwd B=%d D=%d: %.3f ms  (%.1f G densities/s)" % (B, D, ms, B * B * D / ms / 1e6), flush=True)
    zg, mg, lg = (t.clone().requires_grad_(True) for t in (z, mu, lv))
    def tc_step():
        zg.grad = None; mg.grad = None; lg.grad = None
        F.betatc_terms(zg, mg, lg, 0.00025, 1.0, 6.0, 0.37)[0].backward()
    ms = timeit(tc_step, warm=1, it=3)
    print("betatc fwd+bwd B=%d D=%d: %.3f ms" % (B, D, ms), flush=True)
if "swd" in which:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from vaekit import _lib as L
    N, D, S = 1 << 20, 128, 200
    z = 1.5 * torch.randn(N, D, device=dev) + 0.1; p = torch.randn(N, D, device=dev)
    th = torch.randn(S, D, device=dev); th = th / th.norm(dim=1, keepdim=True)
    ms = timeit(lambda: F.swd_sweep(z, p, th, 2.0), warm=1, it=3)
    print("swd N=%d S=%d fwd+grad: %.3f ms  %.1f Msamples/s  (model 28.8 GB -> %.0f GB/s)" % (N, S, ms, N / ms / 1e3, 28.8e9 / ms / 1e6), flush=True)
    L.EVENT_LOG = []
    F.swd_sweep(z, p, th, 2.0)
    torch.cuda.synchronize()
    keys = torch.randn(S, N, device=dev)
    ms = timeit(lambda: F.seg_sort(keys), warm=1, it=3)
    print("seg_sort key-only %dx%d (incl. clone): %.3f ms" % (S, N, ms))
    ms = timeit(lambda: F.seg_sort(keys, want_perm=True), warm=1, it=3)
    print("seg_sort key+perm %dx%d (incl. clone, perm.long()): %.3f ms" % (S, N, ms))
    ms = timeit(lambda: torch.sort(keys, dim=1), warm=1, it=3)
    print("torch.sort %dx%d: %.3f ms" % (S, N, ms))
