"""Parity of the CUDA streaming kernels (K1 reparam+KL, K2 ELBO losses, K3 IW bound),
called through the drop-in model methods / C-ABI, against the golden fixtures made
from the unmodified reference and against the CPU oracle on seeded inputs.
Tolerance: 1e-5 relative (north star), fp32."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import vae_oracle as O  # noqa: E402
from goldens import assert_close, leaf, load  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-5
DEV = "cuda"


@pytest.fixture(scope="module")
def vk():
    import vaekit
    from vaekit import functional as F
    from vaekit.models import vae_models
    vaekit.lib()
    return F, vae_models


def bare(cls, *a, **kw):
    """Model instance without conv stacks on the GPU (only the hot-path methods are exercised)."""
    return cls(*a, **kw)


def test_philox_bits_and_normals(vk):
    F, _ = vk
    raw = F.philox_raw(4096, torch.device(DEV), 1265, 77).cpu().numpy().view(np.uint32)
    assert np.array_equal(raw, O.philox_uint32(1265, 77, 4096))                  # bit-exact integer stream
    big = (1 << 63) + 12345
    raw = F.philox_raw(64, torch.device(DEV), big, (1 << 40) + 3).cpu().numpy().view(np.uint32)
    assert np.array_equal(raw, O.philox_uint32(big, (1 << 40) + 3, 64))
    n = F.philox_normal((100003,), torch.device(DEV), rng=(1265, 8)).cpu().numpy()
    ref = O.philox_normal(1265, 8, 100003)
    assert np.abs(n - ref).max() < 2e-5
    x = F.philox_normal((1 << 22,), torch.device(DEV), rng=(7, 0)).double()
    assert abs(x.mean().item()) < 2e-3 and abs(x.std().item() - 1) < 2e-3
    assert abs((x ** 4).mean().item() - 3) < 3e-2


def test_philox_distribution_and_partition_invariance(vk):
    """SURVEY 8(c) RNG contract: moments and a Kolmogorov-Smirnov distance on 2^26 normals, and bit-reproducibility of a
    fixed (seed, offset) stream however the [N, D] sample is partitioned over 1/2/4/8 ranks (each rank starts the stream
    at its first row: vk_philox_normal_dev with a counter shift, key held on the device)."""
    F, _ = vk
    dev = torch.device(DEV)
    n = 1 << 26
    x = F.philox_normal((n,), dev, rng=(1265, 0))
    xd = x.double()
    m, sd = xd.mean().item(), xd.std().item()
    z = (xd - m) / sd
    assert abs(m) < 5e-4 and abs(sd - 1) < 5e-4
    assert abs((z ** 3).mean().item()) < 2e-3 and abs((z ** 4).mean().item() - 3) < 5e-3
    srt = torch.sort(x[: 1 << 24]).values.double()
    cdf = 0.5 * (1 + torch.erf(srt / 2 ** 0.5))
    emp = (torch.arange(1, srt.numel() + 1, device=dev, dtype=torch.float64)) / srt.numel()
    ks = (cdf - emp).abs().max().item()
    assert ks < 1.63 / (srt.numel() ** 0.5) * 1.5, ks                  # 1 % critical value of the KS statistic, with slack
    # partition invariance
    N, D = 1024, 128
    seed, offset = (1 << 63) + 99, 12345
    full = F.philox_normal((N, D), dev, rng=(seed, offset))
    key = torch.tensor([seed - (1 << 64), offset], dtype=torch.int64, device=dev)
    assert torch.equal(F.philox_normal_dev((N, D), key), full)
    for world in (2, 4, 8):
        rows = N // world
        parts = [F.philox_normal_dev((rows, D), key, quad_shift=r * rows * D // 4) for r in range(world)]
        assert torch.equal(torch.cat(parts), full), world


def test_reparam_matches_golden_with_injected_noise(vk):
    F, _ = vk
    g = load("vanilla", DEV)
    mu, lv = leaf(g["in_mu"]), leaf(g["in_log_var"])
    z, kl, eps = F.reparam_kl(mu, lv, rng=(1265, 0), want_eps=True)
    # hand the kernel's own noise to the oracle (noise injection)
    mu_c, lv_c = leaf(g["in_mu"].cpu()), leaf(g["in_log_var"].cpu())
    z_ref = O.reparameterize(mu_c, lv_c, eps.cpu())
    assert_close(z, z_ref, TOL, "z")
    assert np.abs(eps.cpu().numpy().reshape(-1) - O.philox_normal(1265, 0, eps.numel())).max() < 2e-5
    gz = g["in_gz"]
    (z * gz).sum().backward()
    (z_ref * gz.cpu()).sum().backward()
    assert_close(mu.grad, mu_c.grad, TOL)
    assert_close(lv.grad, lv_c.grad, TOL)
    assert_close(kl, O.kld_rows(g["in_mu"].cpu(), g["in_log_var"].cpu()), TOL, "kl rows")


@pytest.mark.parametrize("shape", [(7, 10), (3, 5, 16), (4, 3, 5, 128), (1, 1), (33, 130)])
def test_reparam_shapes_and_strides(vk, shape):
    F, _ = vk
    gen = torch.Generator().manual_seed(3)
    mu = torch.randn(*shape, generator=gen).to(DEV)
    lv = (0.5 * torch.randn(*shape, generator=gen) - 1).to(DEV)
    mu_l, lv_l = leaf(mu), leaf(lv)
    z, kl, eps = F.reparam_kl(mu_l, lv_l, rng=(5, 4), want_eps=True)
    assert z.shape == mu.shape and z.is_contiguous() and kl.shape == mu.shape[:-1]
    mu_c, lv_c = leaf(mu.cpu()), leaf(lv.cpu())
    z_ref = O.reparameterize(mu_c, lv_c, eps.cpu())
    kl_ref = O.kld_rows(mu_c, lv_c)
    assert_close(z, z_ref, TOL)
    assert_close(kl, kl_ref, TOL)
    gz = torch.randn(*shape, generator=gen)
    gk = torch.randn(*shape[:-1], generator=gen)
    ((z * gz.to(DEV)).sum() + (kl * gk.to(DEV)).sum()).backward()
    ((z_ref * gz).sum() + (kl_ref * gk).sum()).backward()
    assert_close(mu_l.grad, mu_c.grad, TOL)
    assert_close(lv_l.grad, lv_c.grad, TOL)


def test_reparam_expanded_views_and_geometry_independence(vk):
    """IWAE/MIWAE feed stride-0 / permuted views (iwae.py:123-124); the noise must only depend on
    (seed, offset, element index)."""
    F, _ = vk
    gen = torch.Generator().manual_seed(4)
    mu0 = torch.randn(6, 128, generator=gen).to(DEV)
    lv0 = torch.randn(6, 128, generator=gen).to(DEV) * 0.3
    a = F.reparam_kl(mu0.unsqueeze(1).expand(-1, 5, -1), lv0.unsqueeze(1).expand(-1, 5, -1), rng=(9, 16))[0]
    b = F.reparam_kl(mu0.repeat(5, 1, 1).permute(1, 0, 2), lv0.repeat(5, 1, 1).permute(1, 0, 2), rng=(9, 16))[0]
    c = F.reparam_kl(mu0.unsqueeze(1).expand(-1, 5, -1).contiguous(), lv0.unsqueeze(1).expand(-1, 5, -1).contiguous(),
                     rng=(9, 16))[0]
    assert torch.equal(a, b) and torch.equal(a, c)
    # sharding the rows over "ranks" with the matching offset reproduces the same noise
    full = F.reparam_kl(mu0, lv0, rng=(9, 0))[0]
    lo = F.reparam_kl(mu0[:3], lv0[:3], rng=(9, 0))[0]
    hi = F.reparam_kl(mu0[3:], lv0[3:], rng=(9, 3 * 128 // 4))[0]
    assert torch.equal(full, torch.cat([lo, hi]))


def test_vanilla_loss_golden(vk):
    _, models = vk
    g = load("vanilla", DEV)
    m = bare(models["VanillaVAE"], 3, 16)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert d[k].shape == () and d[k].dtype == torch.float32
        assert_close(d[k], g["out_" + k], TOL, k)
    assert not d["Reconstruction_Loss"].requires_grad and not d["KLD"].requires_grad
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)


def test_vanilla_fused_reparam_then_loss(vk):
    """reparameterize() hands its KL rows to loss_function() when the same mu/log_var come back."""
    F, models = vk
    g = load("vanilla", DEV)
    m = bare(models["VanillaVAE"], 3, 16)
    mu, lv, r = leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_recons"])
    z = m.reparameterize(mu, lv)
    assert m._kl._entry is not None
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    assert m._kl._entry is None                     # consumed
    (d["loss"] + (z * g["in_gz"]).sum()).backward()
    assert_close(d["loss"], g["out_loss"], TOL)
    assert_close(r.grad, g["grad_recons"], TOL)
    # mu.grad = dloss/dmu + gz  (dz/dmu = 1)
    assert_close(mu.grad, g["grad_mu"] + g["in_gz"], TOL)


@pytest.mark.parametrize("lt", ["H", "B"])
def test_beta_golden(vk, lt):
    _, models = vk
    g = load("beta_" + lt, DEV)
    kw = dict(beta=10) if lt == "H" else dict(gamma=10.0, max_capacity=25, Capacity_max_iter=10000)
    m = bare(models["BetaVAE"], 3, 16, loss_type=lt, **kw)
    for call in (1, 2, 3):
        r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
        d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
        d["loss"].sum().backward()
        assert tuple(d["loss"].shape) == tuple(g[f"out_loss_{call}"].shape)
        assert_close(d["loss"], g[f"out_loss_{call}"], TOL)
        assert_close(r.grad, g[f"grad_recons_{call}"], TOL)
        assert_close(mu.grad, g[f"grad_mu_{call}"], TOL)
        assert_close(lv.grad, g[f"grad_log_var_{call}"], TOL)
    assert m.num_iter == 3 and type(m).num_iter == 0
    assert d["KLD"].requires_grad and d["Reconstruction_Loss"].requires_grad
    assert_close(d["KLD"], g["out_KLD"], TOL)
    assert_close(d["Reconstruction_Loss"], g["out_Reconstruction_Loss"], TOL)


def test_beta_B_sign_flip_and_side_outputs(vk):
    _, models = vk
    g = load("beta_B_sign", DEV)
    m = bare(models["BetaVAE"], 3, 16, loss_type="B", gamma=10.0, max_capacity=25, Capacity_max_iter=4)
    for call in range(1, 6):
        r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
        d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
        d["loss"].sum().backward()
        assert_close(d["loss"], g[f"out_loss_{call}"], TOL)
        assert_close(mu.grad, g[f"grad_mu_{call}"], TOL)
        assert_close(lv.grad, g[f"grad_log_var_{call}"], TOL)
    # gradients through the non-detached side outputs (beta_vae.py:152)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    (2.0 * d["KLD"] + 3.0 * d["Reconstruction_Loss"]).backward()
    rc, muc, lvc = leaf(g["in_recons"].cpu()), leaf(g["in_mu"].cpu()), leaf(g["in_log_var"].cpu())
    o = O.beta_loss(rc, g["in_input"].cpu(), muc, lvc, g["hp_M_N"], loss_type="B", beta=4, gamma=10.0, C_max=25,
                    C_stop_iter=4, num_iter=6)
    (2.0 * o["KLD"] + 3.0 * o["Reconstruction_Loss"]).backward()
    assert_close(r.grad, rc.grad, TOL); assert_close(mu.grad, muc.grad, TOL); assert_close(lv.grad, lvc.grad, TOL)


def test_beta_undefined_loss_type(vk):
    _, models = vk
    g = load("beta_H", DEV)
    m = bare(models["BetaVAE"], 3, 16, loss_type="X")
    with pytest.raises(ValueError, match="Undefined loss type"):
        m.loss_function(g["in_recons"], g["in_input"], g["in_mu"], g["in_log_var"], M_N=0.005)


def test_logcosh_golden_and_overflow(vk):
    _, models = vk
    g = load("logcosh", DEV)
    m = bare(models["LogCoshVAE"], 3, 16, alpha=g["hp_alpha"], beta=g["hp_beta"])
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)
    # literal form overflows exactly like logcosh_vae.py:146-148
    m = bare(models["LogCoshVAE"], 3, 8, alpha=100.0, beta=1.0)
    r = leaf(torch.full((2, 3, 4, 4), -1.0, device=DEV))
    d = m.loss_function(r, torch.zeros(2, 3, 4, 4, device=DEV), torch.zeros(2, 8, device=DEV),
                        torch.zeros(2, 8, device=DEV), M_N=0.1)
    d["loss"].backward()
    assert torch.isinf(d["loss"]) and torch.isnan(r.grad).all()


def _expand(g, name):
    mu0, lv0 = g["in_mu0"], g["in_log_var0"]
    if name == "miwae":
        return mu0.view(-1, 1, 1, 16).expand(-1, 3, 5, -1), lv0.view(-1, 1, 1, 16).expand(-1, 3, 5, -1)
    return mu0.unsqueeze(1).expand(-1, 5, -1), lv0.unsqueeze(1).expand(-1, 5, -1)


@pytest.mark.parametrize("name", ["iwae", "iwae_w", "miwae"])
def test_iw_golden(vk, name):
    _, models = vk
    g = load(name, DEV)
    m = bare(models["MIWAE"], 3, 16, num_samples=5, num_estimates=3) if name == "miwae" else \
        bare(models["IWAE"], 3, 16, num_samples=5)
    mu, lv = _expand(g, name)
    r, mu, lv = leaf(g["in_recons"]), leaf(mu.contiguous()), leaf(lv.contiguous())
    d = m.loss_function(r, g["in_input"], mu, lv, None, None, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
        assert d[k].requires_grad
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)


@pytest.mark.parametrize("B,G,K,P", [(3, 1, 5, 12288), (2, 3, 5, 12288), (5, 1, 1, 37), (2, 2, 7, 1001), (1, 1, 3, 50000)])
def test_iw_shapes_vs_oracle(vk, B, G, K, P):
    F, _ = vk
    gen = torch.Generator().manual_seed(B * 100 + K)
    shape = (B, G, K) if G > 1 else (B, K)
    r = torch.tanh(torch.randn(*shape, 1, 1, P, generator=gen))      # [B,(G,)K,C,H,W] with C=H=1
    x = torch.rand(B, 1, 1, P, generator=gen)
    mu = torch.randn(*shape, 8, generator=gen)
    lv = 0.5 * torch.randn(*shape, 8, generator=gen)
    rc, muc, lvc = leaf(r), leaf(mu), leaf(lv)
    o = O.iw_bound_loss(rc, x, muc, lvc, 0.01)
    (o["loss"] + 0.3 * o["Reconstruction_Loss"] - 0.2 * o["KLD"]).backward()
    rg, mug, lvg = leaf(r.to(DEV)), leaf(mu.to(DEV)), leaf(lv.to(DEV))
    loss, recon, _, negkld = F.iw_bound(rg, x.to(DEV), F.kl_rows(mug, lvg), 0.01)
    (loss + 0.3 * recon - 0.2 * negkld).backward()
    assert_close(loss, o["loss"], TOL); assert_close(recon, o["Reconstruction_Loss"], TOL)
    assert_close(negkld, o["KLD"], TOL)
    assert_close(rg.grad, rc.grad, TOL); assert_close(mug.grad, muc.grad, TOL); assert_close(lvg.grad, lvc.grad, TOL)


@pytest.mark.parametrize("mode", ["mse_mean", "mse_sum", "mse_l1", "logcosh"])
@pytest.mark.parametrize("numel", [1, 5, 4099, 1 << 20])
def test_elbo_recon_modes_vs_torch(vk, mode, numel):
    F, _ = vk
    gen = torch.Generator().manual_seed(numel)
    r = torch.tanh(torch.randn(numel, generator=gen)); x = torch.rand(numel, generator=gen)
    rc = leaf(r)
    tf = torch.nn.functional
    if mode == "mse_mean":
        ref = tf.mse_loss(rc, x); kw = dict(recon_mode=F.RECON_MSE_MEAN)
    elif mode == "mse_sum":
        ref = tf.mse_loss(rc, x, reduction="sum"); kw = dict(recon_mode=F.RECON_MSE_SUM)
    elif mode == "mse_l1":
        ref = tf.mse_loss(rc, x) + tf.l1_loss(rc, x); kw = dict(recon_mode=F.RECON_MSE_L1_MEAN)
    else:
        ref = O.logcosh_loss(rc, x, torch.zeros(1, 2), torch.zeros(1, 2), 0.0, alpha=10.0, beta=1.0)["Reconstruction_Loss"]
        kw = dict(recon_mode=F.RECON_LOGCOSH, alpha=10.0)
    ref.backward()
    rg = leaf(r.to(DEV))
    loss, recon, _, _ = F.elbo(rg, x.to(DEV), None, **kw)
    loss.backward()
    assert_close(recon, ref, TOL); assert_close(loss, ref, TOL)
    assert_close(rg.grad, rc.grad, TOL)


def test_elbo_unaligned_views(vk):
    F, _ = vk
    gen = torch.Generator().manual_seed(11)
    base_r = torch.randn(4100, generator=gen).to(DEV); base_x = torch.randn(4100, generator=gen).to(DEV)
    r, x = base_r[1:4098], base_x[3:4100]                      # 4-byte aligned only
    rg = leaf(r)
    loss = F.elbo(rg, x, None)[0]
    loss.backward()
    rc = leaf(r.cpu())
    ref = torch.nn.functional.mse_loss(rc, x.cpu()); ref.backward()
    assert_close(loss, ref, TOL); assert_close(rg.grad, rc.grad, TOL)


def test_rejects_cpu_and_wrong_dtype(vk):
    F, _ = vk
    with pytest.raises(RuntimeError, match="CUDA"):
        F.elbo(torch.zeros(4), torch.zeros(4), None)
    with pytest.raises(RuntimeError, match="float32"):
        F.elbo(torch.zeros(4, device=DEV, dtype=torch.float64), torch.zeros(4, device=DEV, dtype=torch.float64), None)
    with pytest.raises(RuntimeError, match="shape"):
        F.elbo(torch.zeros(4, device=DEV), torch.zeros(5, device=DEV), None)


def _iw_fp64(recons, x, mu0, lv0, M_N, G, K):
    """iwae.py:145-158 / miwae.py:148-162 evaluated in float64 with torch ON THE GPU (torch is only the checker): the
    input is broadcast instead of repeated, everything else is the reference's sequence of ops."""
    r = recons.double().requires_grad_(True)
    m = mu0.double().requires_grad_(True); v = lv0.double().requires_grad_(True)
    shp = (r.shape[0],) + (1,) * (r.dim() - 4) + tuple(x.shape[1:])
    lp = ((r - x.double().view(shp)) ** 2).flatten(r.dim() - 3).mean(-1)            # [B,(M,)K]
    ms = (mu0.shape[0],) + (1,) * (r.dim() - 4) + (mu0.shape[1],)
    kld = -0.5 * torch.sum(1 + v.view(ms) - m.view(ms) ** 2 - v.view(ms).exp(), dim=-1).expand_as(lp)
    lw = lp + M_N * kld
    w = torch.softmax(lw, dim=-1)
    loss = torch.sum(w * lw, dim=-1).mean()
    loss.backward()
    return loss, lp.mean(), -kld.mean(), r.grad, m.grad, v.grad


@pytest.mark.parametrize("G", [1, 3])
def test_iw_bound_at_benchmark_size_vs_fp64(vk, G):
    """BASELINE configs[1] at the size bench.py times: IWAE K=5 (and MIWAE M=3) at B=4096, D=128, 3x64x64.  The loss, its
    two side outputs and all three gradients (recons, mu, log_var through the stride-0 expand) against the float64
    restatement of iwae.py:145-158 / miwae.py:148-162 on the same tensors, 1e-5 relative."""
    F, vae_models = vk
    B, K, D, M_N = 4096, 5, 128, 0.00025
    gen = torch.Generator(device=DEV).manual_seed(1265)
    ss = (B, G, K) if G > 1 else (B, K)
    r = torch.tanh(torch.randn(*ss, 3, 64, 64, device=DEV, generator=gen))
    x = torch.rand(B, 3, 64, 64, device=DEV, generator=gen)
    mu0 = torch.randn(B, D, device=DEV, generator=gen); lv0 = 0.5 * torch.randn(B, D, device=DEV, generator=gen) - 1
    m = (vae_models["MIWAE"](3, D, num_samples=K, num_estimates=G) if G > 1 else vae_models["IWAE"](3, D, num_samples=K))
    rg, mg, vg = leaf(r), leaf(mu0), leaf(lv0)
    es = (-1, G, K, -1) if G > 1 else (-1, K, -1)
    mu = mg.view(B, *([1] * (len(ss) - 1)), D).expand(*es); lv = vg.view(B, *([1] * (len(ss) - 1)), D).expand(*es)
    d = m.loss_function(rg, x, mu, lv, None, None, M_N=M_N)
    d["loss"].backward()
    # float64 reference in row chunks of the batch (3 GB of recons in double does not fit twice): the loss is a mean over
    # rows, so chunk gradients are scaled by chunk/B and the chunk losses averaged
    loss = rl = kl = 0.0
    g_r = torch.empty_like(r); g_m = torch.empty_like(mu0); g_v = torch.empty_like(lv0)
    CH = 512
    for b0 in range(0, B, CH):
        sl = slice(b0, b0 + CH)
        l, a, k, gr, gm, gv = _iw_fp64(r[sl], x[sl], mu0[sl], lv0[sl], M_N, G, K)
        loss += l.item() * CH / B; rl += a.item() * CH / B; kl += k.item() * CH / B
        g_r[sl] = (gr * (CH / B)).float(); g_m[sl] = (gm * (CH / B)).float(); g_v[sl] = (gv * (CH / B)).float()
    assert abs(d["loss"].item() - loss) <= TOL * abs(loss)
    assert abs(d["Reconstruction_Loss"].item() - rl) <= TOL * abs(rl)
    assert abs(d["KLD"].item() - kl) <= TOL * abs(kl)
    assert_close(rg.grad, g_r, TOL, "g_recons"); assert_close(mg.grad, g_m, TOL, "g_mu"); assert_close(vg.grad, g_v, TOL, "g_log_var")
    # and per row: the recons gradient of every sample row, relative to that row's own maximum (the global max-norm above
    # would hide a wrong low-weight sample)
    num = (rg.grad - g_r).flatten(len(ss)).abs().amax(-1); den = g_r.flatten(len(ss)).abs().amax(-1).clamp_min(1e-30)
    assert (num / den).max().item() <= 2e-5


def _load_forward_fixture(name, cls, dev):
    fx = load(name)
    D, K = int(fx["hp_D"]), int(fx["hp_K"])
    kw = dict(num_samples=K)
    if "hp_M" in fx:
        kw["num_estimates"] = int(fx["hp_M"])
    m = cls(3, D, hidden_dims=[2, 4, 4, 8, 512], **kw).eval()
    m.load_state_dict({k[2:]: v for k, v in fx.items() if k.startswith("w_")}, strict=True)
    return fx, m.to(dev)


@pytest.mark.parametrize("name", ["iwae_forward", "miwae_forward"])
def test_iw_forward_contract_and_fixture(vk, name):
    """IWAE.forward / MIWAE.forward / generate (iwae.py:121-127,188, miwae.py:124-130,192) against the fixture made from
    the reference classes with their own weights: the 6-element list, shapes, a dense [B,(M,)S,D] z,
    eps = (z - mu) / log_var, and -- with the fixture's noise handed in -- every returned tensor."""
    F, vae_models = vk
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
    try:
        fx, m = _load_forward_fixture(name, vae_models["MIWAE" if "hp_M" in load(name) else "IWAE"], DEV)
        x = fx["in_input"].to(DEV)
        B, D, K = x.shape[0], int(fx["hp_D"]), int(fx["hp_K"])
        ss = (B, int(fx["hp_M"]), K) if "hp_M" in fx else (B, K)
        with torch.no_grad():
            torch.manual_seed(5)
            out = m(x, labels=None)                              # forward must swallow labels= (experiment.py:36)
        assert isinstance(out, list) and len(out) == 6 and out[1] is x
        rec, _, mu, lv, z, eps = out
        assert tuple(rec.shape) == ss + (3, 64, 64) and all(tuple(t.shape) == ss + (D,) for t in (mu, lv, z, eps))
        assert z.is_contiguous() and rec.is_contiguous()
        assert torch.equal(eps, (z - mu) / lv)
        assert_close(mu, fx["out_mu"].to(DEV), 1e-4, "mu"); assert_close(lv, fx["out_log_var"].to(DEV), 1e-4, "log_var")
        # the kernel's own noise: independent per sample (not one draw broadcast over the sample dims), and the same
        # stream F.reparam_kl draws for the same generator state
        n = (z - mu) * torch.exp(-0.5 * lv)
        assert (n.flatten(1)[:, :D] - n.flatten(1)[:, D:2 * D]).abs().min().item() > 0
        torch.manual_seed(5)
        z2, _, n2 = F.reparam_kl(mu, lv, want_eps=True)
        assert torch.equal(z2, z) and (n2 - n).abs().max().item() < 1e-5
        # the fixture's noise handed in: every tensor of the reference's forward
        inj = fx["noise_eps"].to(DEV)
        m.reparameterize = lambda a, b: inj * torch.exp(0.5 * b) + a
        with torch.no_grad():
            out = m(x)
            gen = m.generate(x)
        for got, key in zip((out[0], out[2], out[3], out[4], out[5]), ("recons", "mu", "log_var", "z", "eps")):
            assert_close(got, fx["out_" + key].to(DEV), 1e-4, key)
        assert tuple(gen.shape) == (B, 3, 64, 64)
        assert_close(gen, fx["out_generate"].to(DEV), 1e-4, "generate")
        del m.reparameterize
        with torch.no_grad():
            s = m.sample(5, torch.device(DEV))
        assert tuple(s.shape) == (5, 3, 64, 64)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32


@pytest.mark.parametrize("reduction", ["mean", "sum"])
def test_bce_reconstruction_mode(vk, reduction):
    """North star kernel (1) "MSE/BCE": F.binary_cross_entropy semantics incl. the clamp of both logs at -100 (recons
    exactly 0 / 1) and the 1e-12 floor of the backward's denominator; oracle = torch's own op (the reference has no BCE)."""
    F, vae_models = vk
    g = torch.Generator().manual_seed(77)
    r = torch.sigmoid(2 * torch.randn(6, 3, 16, 16, generator=g)); x = torch.rand(6, 3, 16, 16, generator=g)
    r.view(-1)[:4] = torch.tensor([0.0, 1.0, 1e-30, 1 - 1e-7])                    # clamp / floor cases
    x.view(-1)[:2] = torch.tensor([1.0, 0.0])                                     # log(0) with a non-zero weight
    mu = torch.randn(6, 16, generator=g); lv = 0.5 * torch.randn(6, 16, generator=g) - 1
    rc, mc, vc = leaf(r), leaf(mu), leaf(lv)
    ref = O.bce_vanilla_loss(rc, x, mc, vc, 0.005, reduction=reduction)
    ref["loss"].backward()
    rg, mg, vg = leaf(r.to(DEV)), leaf(mu.to(DEV)), leaf(lv.to(DEV))
    mode = F.RECON_BCE_MEAN if reduction == "mean" else F.RECON_BCE_SUM
    loss, rec, _, negkld = F.elbo(rg, x.to(DEV), F.kl_rows(mg, vg), recon_mode=mode, combine_mode=F.COMBINE_LINEAR, w=0.005)
    loss.backward()
    assert_close(loss, ref["loss"], TOL); assert_close(rec, ref["Reconstruction_Loss"], TOL); assert_close(negkld, ref["KLD"], TOL)
    # the floor cases produce gradients of 1e12 magnitude: compare them element-wise relative, the rest in max norm
    big = rc.grad.abs() > 1e3
    assert torch.allclose(rg.grad.cpu()[big], rc.grad[big], rtol=1e-5)
    assert_close(rg.grad.cpu()[~big], rc.grad[~big], TOL)
    assert_close(mg.grad, mc.grad, TOL); assert_close(vg.grad, vc.grad, TOL)
    if reduction == "mean":                                                       # through the drop-in model
        m = vae_models["VanillaVAE"](3, 16, recons_loss="bce")
        d = m.loss_function(r.to(DEV), x.to(DEV), mu.to(DEV), lv.to(DEV), M_N=0.005)
        for k in d:
            assert_close(d[k], ref[k], TOL, k)
        with pytest.raises(ValueError):
            vae_models["VanillaVAE"](3, 16, recons_loss="l7")


def test_device_rng_key_and_graph_replay(vk):
    """(seed, offset) read from device memory (vk_reparam_kl_fwd_dev / _bwd_dev).  Same stream as the
    by-value entry points; a CUDA graph that captured reparameterize + loss + backward draws NEW noise on every replay
    because the key advances on the device, and its backward regenerates the noise of its own forward."""
    F, vae_models = vk
    dev = torch.device(DEV)
    g = torch.Generator(device=DEV).manual_seed(3)
    B, D = 64, 128
    mu = torch.randn(B, D, device=DEV, generator=g); lv = 0.5 * torch.randn(B, D, device=DEV, generator=g) - 1
    key = torch.tensor([1265, 40], dtype=torch.int64, device=DEV)
    z_val, _, e_val = F.reparam_kl(mu, lv, rng=(1265, 40), want_eps=True)
    z_dev, _, e_dev = F.reparam_kl(mu, lv, rng=key.clone(), want_eps=True)
    assert torch.equal(z_val, z_dev) and torch.equal(e_val, e_dev)
    # graph: static inputs, the model's own reparameterize + loss + backward inside device_rng
    m = vae_models["VanillaVAE"](3, D)
    r = torch.tanh(torch.randn(B, 3, 64, 64, device=DEV, generator=g)); x = torch.rand(B, 3, 64, 64, device=DEV, generator=g)
    mus, lvs = leaf(mu), leaf(lv)
    gz = torch.randn(B, D, device=DEV, generator=g)
    key = F.make_rng_key(dev)
    key0 = key.clone()

    def step():
        mus.grad = lvs.grad = None
        z = m.reparameterize(mus, lvs)
        d = m.loss_function(r, x, mus, lvs, M_N=0.005)
        torch.autograd.backward([d["loss"], z], [None, gz])
        return z, d["loss"]

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), F.device_rng(key):
        for _ in range(2):
            step()
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    mus.grad = lvs.grad = None
    with F.device_rng(key):
        with torch.cuda.graph(graph, stream=s):             # the warmed-up stream: its workspace already exists
            z_out, loss_out = step()
    zs, gls = [], []
    for _ in range(3):
        k_before = key.clone()
        graph.replay()
        torch.cuda.synchronize()
        zs.append(z_out.clone()); gls.append(lvs.grad.clone())
        # this replay's forward AND backward used the key as it stood before the replay
        z_chk, _, eps = F.reparam_kl(mu, lv, rng=k_before, want_eps=True)
        assert torch.equal(z_chk, z_out)
        want_glv = gz * eps * 0.5 * torch.exp(0.5 * lv) - 0.5 * (0.005 / B) * (1 - torch.exp(lv))
        assert_close(lvs.grad, want_glv, TOL)
        assert int(key[1]) == int(k_before[1]) + B * D // 4 and int(key[0]) == int(key0[0])
    assert not torch.equal(zs[0], zs[1]) and not torch.equal(zs[1], zs[2])
