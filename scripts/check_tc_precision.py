"""Who is closer to float64: the CUDA BetaTC kernels or the fp32 oracle?  (the parity tests compare against fp32)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
import vae_oracle as O
from vaekit import functional as F

def rel(a, b): return ((a.double().cpu() - b.double().cpu()).abs().max() / b.double().abs().max()).item()
for B, D in ((130, 128), (1024, 128), (1030, 40)):
    gen = torch.Generator().manual_seed(B + D)
    mu = torch.randn(B, D, generator=gen); lv = 0.3 * torch.randn(B, D, generator=gen)
    z = mu + torch.randn(B, D, generator=gen) * torch.exp(0.5 * lv)
    rz = torch.zeros(B, 1, 1, 1)
    res = {}
    for name, dt in (("f32", torch.float32), ("f64", torch.float64)):
        zc, mc, lc = (t.to(dt).clone().requires_grad_(True) for t in (z, mu, lv))
        ref = O.betatc_loss(rz.to(dt), rz.to(dt), mc, lc, zc, 0.00025, alpha=1.0, beta=6.0, gamma=1.0, anneal_rate=0.37)
        ref["loss"].backward()
        res[name] = (ref["loss"].detach(), zc.grad, mc.grad, lc.grad)
    zg, mg, lg = (t.cuda().clone().requires_grad_(True) for t in (z, mu, lv))
    latent, terms = F.betatc_terms(zg, mg, lg, 0.00025, 1.0, 6.0, 0.37)
    latent.backward()
    ours = (latent.detach(), zg.grad, mg.grad, lg.grad)
    for k, n in enumerate(("loss", "g_z", "g_mu", "g_lv")):
        print("B=%d D=%d %-5s ours-vs-f64 %.2e   oracle32-vs-f64 %.2e   ours-vs-oracle32 %.2e" %
              (B, D, n, rel(ours[k], res["f64"][k]), rel(res["f32"][k], res["f64"][k]), rel(ours[k], res["f32"][k])), flush=True)
