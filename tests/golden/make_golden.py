"""Generate tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference/models, imported read-only) on small seeded inputs.

Run in the build container only:  python tests/golden/make_golden.py
Array naming: in_* inputs, noise_* injected randn_like draws, hp_* scalar
hyper-parameters, out_* loss-dict entries / returned tensors, grad_* gradients
of out_loss (upstream grad = 1) w.r.t. the matching in_* array.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from refload import injected_noise, load_reference_models  # noqa: E402

ref = load_reference_models()


def leaf(t):
    return t.clone().detach().requires_grad_(True)


def save(name, **arrs):
    out = {}
    for k, v in arrs.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        out[k] = np.asarray(v)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items()})


def images(g, b, hw=16, kind="tanh"):
    r = torch.tanh(torch.randn(b, 3, hw, hw, generator=g))
    x = torch.rand(b, 3, hw, hw, generator=g)
    return r, x


def gaussians(g, *shape):
    return torch.randn(*shape, generator=g), 0.5 * torch.randn(*shape, generator=g) - 1.0


def gen_vanilla():
    g = torch.Generator().manual_seed(1265)
    m = ref.VanillaVAE(3, 16)
    r, x = images(g, 6)
    mu, lv = gaussians(g, 6, 16)
    eps = torch.randn(6, 16, generator=g)
    gz = torch.randn(6, 16, generator=g)
    mu1, lv1 = leaf(mu), leaf(lv)
    with injected_noise(eps):
        z = m.reparameterize(mu1, lv1)
    z.backward(gz)
    rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
    d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
    d["loss"].backward()
    save("vanilla", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, noise_eps=eps, in_gz=gz, hp_M_N=0.005,
         out_z=z, grad_z_mu=mu1.grad, grad_z_log_var=lv1.grad,
         out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)


def gen_beta():
    g = torch.Generator().manual_seed(1266)
    r, x = images(g, 5)
    mu, lv = gaussians(g, 5, 16)
    for lt, kw in (("H", dict(beta=10)), ("B", dict(gamma=10.0, max_capacity=25, Capacity_max_iter=10000))):
        m = ref.BetaVAE(3, 16, loss_type=lt, **kw)
        outs = {}
        for call in (1, 2, 3):      # num_iter increments on every call (beta_vae.py:132)
            rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
            d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
            d["loss"].sum().backward()
            outs[f"out_loss_{call}"] = d["loss"]
            outs[f"grad_recons_{call}"] = rl.grad
            outs[f"grad_mu_{call}"] = mul.grad
            outs[f"grad_log_var_{call}"] = lvl.grad
        save("beta_" + lt, in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005,
             out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"], hp_num_iter=m.num_iter, **outs)
    # large-KLD variant so that (kld - C) changes sign between calls
    m = ref.BetaVAE(3, 16, loss_type="B", gamma=10.0, max_capacity=25, Capacity_max_iter=4)
    mu2 = 0.9 * mu
    outs = {}
    for call in (1, 2, 3, 4, 5):
        rl, mul, lvl = leaf(r), leaf(mu2), leaf(lv)
        d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
        d["loss"].sum().backward()
        outs[f"out_loss_{call}"] = d["loss"]
        outs[f"grad_mu_{call}"] = mul.grad
        outs[f"grad_log_var_{call}"] = lvl.grad
    save("beta_B_sign", in_recons=r, in_input=x, in_mu=mu2, in_log_var=lv, hp_M_N=0.005, out_KLD=d["KLD"], **outs)


def gen_logcosh():
    g = torch.Generator().manual_seed(1267)
    m = ref.LogCoshVAE(3, 16, alpha=10.0, beta=1.0)
    r, x = images(g, 6)
    mu, lv = gaussians(g, 6, 16)
    rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
    d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
    d["loss"].backward()
    save("logcosh", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005, hp_alpha=10.0, hp_beta=1.0,
         out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)


def gen_iw():
    g = torch.Generator().manual_seed(1268)
    B, D, K, M = 4, 16, 5, 3
    x = torch.rand(B, 3, 16, 16, generator=g)
    mu0, lv0 = gaussians(g, B, D)
    # IWAE
    m = ref.IWAE(3, D, num_samples=K)
    mu = mu0.repeat(K, 1, 1).permute(1, 0, 2)
    lv = lv0.repeat(K, 1, 1).permute(1, 0, 2)
    eps = torch.randn(B, K, D, generator=g)
    with injected_noise(eps):
        z = m.reparameterize(mu, lv)
    r = torch.tanh(torch.randn(B, K, 3, 16, 16, generator=g))
    rl, mul, lvl = leaf(r), leaf(mu.contiguous()), leaf(lv.contiguous())
    d = m.loss_function(rl, x, mul, lvl, z, None, M_N=0.00025)
    d["loss"].backward()
    save("iwae", in_recons=r, in_input=x, in_mu0=mu0, in_log_var0=lv0, noise_eps=eps, hp_M_N=0.00025,
         out_z=z, out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)
    # same with a weight large enough that the KL term matters in the softmax
    rl, mul, lvl = leaf(r), leaf(mu.contiguous()), leaf(lv.contiguous())
    d = m.loss_function(rl, x, mul, lvl, z, None, M_N=0.05)
    d["loss"].backward()
    save("iwae_w", in_recons=r, in_input=x, in_mu0=mu0, in_log_var0=lv0, hp_M_N=0.05,
         out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)
    # MIWAE
    m = ref.MIWAE(3, D, num_samples=K, num_estimates=M)
    mu = mu0.repeat(M, K, 1, 1).permute(2, 0, 1, 3)
    lv = lv0.repeat(M, K, 1, 1).permute(2, 0, 1, 3)
    eps = torch.randn(B, M, K, D, generator=g)
    with injected_noise(eps):
        z = m.reparameterize(mu, lv)
    r = torch.tanh(torch.randn(B, M, K, 3, 16, 16, generator=g))
    rl, mul, lvl = leaf(r), leaf(mu.contiguous()), leaf(lv.contiguous())
    d = m.loss_function(rl, x, mul, lvl, z, None, M_N=0.00025)
    d["loss"].backward()
    save("miwae", in_recons=r, in_input=x, in_mu0=mu0, in_log_var0=lv0, noise_eps=eps, hp_M_N=0.00025,
         out_z=z, out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)


def gen_wae():
    g = torch.Generator().manual_seed(1269)
    for kt, rw, n, dim in (("rbf", 5000, 40, 16), ("imq", 100, 40, 16), ("rbf", 5000, 96, 128), ("imq", 100, 96, 128)):
        m = ref.WAE_MMD(3, dim, reg_weight=rw, kernel_type=kt)
        r, x = images(g, n, hw=8)
        z = torch.randn(n, dim, generator=g) + 0.25
        pz = torch.randn(n, dim, generator=g)
        rl, zl = leaf(r), leaf(z)
        with injected_noise(pz):
            d = m.loss_function(rl, x, zl)
        d["loss"].backward()
        with torch.no_grad():
            pp = m.compute_kernel(pz, pz).mean()
            zz = m.compute_kernel(z, z).mean()
            pzt = m.compute_kernel(pz, z).mean()
        save(f"wae_{kt}_d{dim}", in_recons=r, in_input=x, in_z=z, noise_prior=pz, hp_reg_weight=rw, hp_z_var=2.0,
             out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_MMD=d["MMD"],
             out_term_pp=pp, out_term_zz=zz, out_term_pz=pzt, grad_recons=rl.grad, grad_z=zl.grad)


def gen_betatc():
    g = torch.Generator().manual_seed(1270)
    for b, dim in ((16, 10), (48, 32)):
        m = ref.BetaTCVAE(3, dim, anneal_steps=100, alpha=1.0, beta=6.0, gamma=1.0)
        m.train()
        r, x = images(g, b, hw=8)
        mu, lv = torch.randn(b, dim, generator=g), 0.3 * torch.randn(b, dim, generator=g)
        eps = torch.randn(b, dim, generator=g)
        rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
        with injected_noise(eps):
            z = m.reparameterize(mul, lvl)
        d = m.loss_function(rl, x, mul, lvl, z, M_N=0.00025)
        d["loss"].backward()
        # z treated as an independent leaf as well (loss_function called on given tensors)
        rl2, mul2, lvl2, zl2 = leaf(r), leaf(mu), leaf(lv), leaf(z)
        m.eval()
        d2 = m.loss_function(rl2, x, mul2, lvl2, zl2, M_N=0.00025)
        d2["loss"].backward()
        save(f"betatc_b{b}_d{dim}", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, noise_eps=eps,
             hp_M_N=0.00025, hp_anneal_train=1 / 100, out_z=z,
             out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
             out_TC_Loss=d["TC_Loss"], out_MI_Loss=d["MI_Loss"],
             grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad,
             out_loss_eval=d2["loss"], grad_eval_mu=mul2.grad, grad_eval_log_var=lvl2.grad, grad_eval_z=zl2.grad)


def gen_swae():
    g = torch.Generator().manual_seed(1271)
    for n, dim, s in ((64, 16, 20), (256, 128, 50)):
        m = ref.SWAE(3, dim, reg_weight=100, wasserstein_deg=2.0, num_projections=s)
        r, x = images(g, n, hw=8)
        z = 1.5 * torch.randn(n, dim, generator=g) + 0.1
        pz = torch.randn(n, dim, generator=g)
        torch.manual_seed(1265)
        proj = m.get_random_projections(dim, s)           # what the next call will draw
        torch.manual_seed(1265)
        rl, zl = leaf(r), leaf(z)
        with injected_noise(pz):
            d = m.loss_function(rl, x, zl)
        d["loss"].backward()
        save(f"swae_n{n}_d{dim}", in_recons=r, in_input=x, in_z=z, noise_prior=pz, in_proj=proj,
             hp_reg_weight=100, hp_p=2.0, hp_seed=1265,
             out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_SWD=d["SWD"],
             grad_recons=rl.grad, grad_z=zl.grad)


def gen_vq():
    g = torch.Generator().manual_seed(1272)
    for tag, k, dim, shape, init in (("default", 32, 8, (2, 8, 4, 4), "default"), ("normal", 512, 64, (2, 64, 8, 8), "normal")):
        q = ref.VectorQuantizer(k, dim, 0.25)
        with torch.no_grad():
            if init == "normal":
                q.embedding.weight.copy_(torch.randn(k, dim, generator=g))
            else:
                q.embedding.weight.copy_((torch.rand(k, dim, generator=g) * 2 - 1) / k)
        lat = torch.nn.functional.leaky_relu(torch.randn(*shape, generator=g))
        gq = 1e-3 * torch.randn(*shape, generator=g)
        ll = leaf(lat)
        out, vq_loss = q(ll)
        (vq_loss + (out * gq).sum()).backward()
        flat = lat.permute(0, 2, 3, 1).reshape(-1, dim)
        w = q.embedding.weight.detach()
        dist = (flat ** 2).sum(1, keepdim=True) + (w ** 2).sum(1) - 2 * flat @ w.t()
        idx = dist.argmin(1)
        save(f"vq_{tag}", in_latents=lat, in_codebook=w, in_gq=gq, hp_beta=0.25,
             out_quantized=out, out_vq_loss=vq_loss, out_indices=idx,
             grad_latents=ll.grad, grad_codebook=q.embedding.weight.grad)


def gen_info():
    """info_vae.py:117-134 with z chained to (mu, log_var) through reparameterize, as forward() does."""
    g = torch.Generator().manual_seed(1272)
    for kt, n, dim, kw in (("imq", 40, 16, dict(alpha=-9.0, beta=10.5, reg_weight=110)),
                           ("rbf", 96, 128, dict(alpha=-0.5, beta=5.0, reg_weight=100))):
        m = ref.InfoVAE(3, dim, kernel_type=kt, **kw)
        r, x = images(g, n, hw=8)
        mu, lv = gaussians(g, n, dim)
        eps = torch.randn(n, dim, generator=g)
        pz = torch.randn(n, dim, generator=g)
        rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
        with injected_noise(eps, pz):
            z = m.reparameterize(mul, lvl)
            d = m.loss_function(rl, x, z, mul, lvl, M_N=0.005)
        d["loss"].backward()
        with torch.no_grad():
            pp = m.compute_kernel(pz, pz).mean()
            zz = m.compute_kernel(z, z).mean()
            pzt = m.compute_kernel(pz, z).mean()
        save(f"info_{kt}_d{dim}", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, noise_eps=eps, noise_prior=pz,
             hp_M_N=0.005, hp_alpha=kw["alpha"], hp_beta=kw["beta"], hp_reg_weight=kw["reg_weight"], hp_z_var=2.0,
             out_z=z, out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_MMD=d["MMD"],
             out_KLD=d["KLD"], out_term_pp=pp, out_term_zz=zz, out_term_pz=pzt,
             grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)


def gen_next_models():
    """ConditionalVAE, DIPVAE, FactorVAE (VAE branch) loss functions of the unmodified reference."""
    g = torch.Generator().manual_seed(77)
    r, x = images(g, 6)
    mu, lv = gaussians(g, 6, 16)
    # ConditionalVAE
    m = ref.ConditionalVAE(3, 10, 16)
    rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
    d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
    d["loss"].backward()
    save("cvae", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005, out_loss=d["loss"],
         out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"], grad_recons=rl.grad, grad_mu=mul.grad,
         grad_log_var=lvl.grad)
    # DIPVAE
    m = ref.DIPVAE(3, 16, lambda_diag=10., lambda_offdiag=5.)
    rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
    d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
    d["loss"].backward()
    save("dipvae", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005, out_loss=d["loss"],
         out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"], out_DIP_Loss=d["DIP_Loss"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad)
    # TwoStageVAE (first-stage loss)
    m = ref.TwoStageVAE(3, 16)
    rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
    d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
    d["loss"].backward()
    save("twostage", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005, out_loss=d["loss"],
         out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"], grad_recons=rl.grad, grad_mu=mul.grad,
         grad_log_var=lvl.grad)
    # FactorVAE, optimizer_idx = 0; the discriminator is replaced by a fixed linear map so the fixture stays small
    m = ref.FactorVAE(3, 16, gamma=40.)
    W = 0.3 * torch.randn(16, 2, generator=g)
    lin = torch.nn.Linear(16, 2, bias=False)
    with torch.no_grad():
        lin.weight.copy_(W.t())
    m.discriminator = lin
    z = torch.randn(6, 16, generator=g)
    rl, mul, lvl, zl = leaf(r), leaf(mu), leaf(lv), leaf(z)
    d = m.loss_function(rl, x, mul, lvl, zl, M_N=0.005, optimizer_idx=0)
    d["loss"].backward()
    save("factorvae", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, in_z=z, in_W=W, hp_M_N=0.005, out_loss=d["loss"],
         out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"], out_VAE_TC_Loss=d["VAE_TC_Loss"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad, grad_z=zl.grad)


def gen_cat():
    """CategoricalVAE reparameterize (Gumbel-softmax, torch.rand_like patched to return the fixture's
    uniforms) and loss_function of the unmodified reference."""
    g = torch.Generator().manual_seed(404)
    m = ref.CategoricalVAE(3, 6, categorical_dim=40, temperature=0.5, alpha=30.)
    r, x = images(g, 5)
    q = 1.5 * torch.randn(5, 6, 40, generator=g)
    u = torch.rand(5, 6, 40, generator=g)
    gs = torch.randn(5, 240, generator=g)
    ql = leaf(q)
    orig = torch.rand_like
    torch.rand_like = lambda t, **kw: u.to(t.dtype)
    try:
        s = m.reparameterize(ql)
    finally:
        torch.rand_like = orig
    s.backward(gs)
    rl, q2 = leaf(r), leaf(q)
    m.eval()                                           # no annealing: the temperature stays at its constructor value
    d = m.loss_function(rl, x, q2, M_N=0.005, batch_idx=3)
    d["loss"].backward()
    save("catvae", in_recons=r, in_input=x, in_q=q, noise_u=u, in_gs=gs, hp_M_N=0.005, hp_temp=0.5, hp_alpha=30.,
         out_s=s, grad_s_q=ql.grad, out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"],
         out_KLD=d["KLD"], grad_recons=rl.grad, grad_q=q2.grad)


def gen_joint():
    """JointVAE reparameterize (randn_like / rand_like patched) and two training-mode loss calls
    (num_iter 1 then 2) of the unmodified reference."""
    g = torch.Generator().manual_seed(505)
    kw = dict(latent_num_iter=50, categorical_num_iter=20, latent_max_capacity=25., categorical_max_capacity=25.)
    m = ref.JointVAE(3, 12, 10, **kw)
    r, x = images(g, 6)
    mu, lv = gaussians(g, 6, 12)
    q = 1.5 * torch.randn(6, 10, generator=g)
    e, u = torch.randn(6, 12, generator=g), torch.rand(6, 10, generator=g)
    gz = torch.randn(6, 22, generator=g)
    mul, lvl, ql = leaf(mu), leaf(lv), leaf(q)
    orig = torch.rand_like
    torch.rand_like = lambda t, **k2: u.to(t.dtype)
    try:
        with injected_noise(e):
            z = m.reparameterize(mul, lvl, ql)
    finally:
        torch.rand_like = orig
    z.backward(gz)
    out = dict(in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, in_q=q, noise_eps=e, noise_u=u, in_gz=gz, hp_M_N=0.005,
               out_z=z, grad_z_mu=mul.grad, grad_z_log_var=lvl.grad, grad_z_q=ql.grad)
    m.train()
    for call in (1, 2):
        rl, mul, lvl, ql = leaf(r), leaf(mu), leaf(lv), leaf(q)
        d = m.loss_function(rl, x, ql, mul, lvl, M_N=0.005, batch_idx=7)
        d["loss"].backward()
        out.update({"out_loss_%d" % call: d["loss"], "out_Capacity_Loss_%d" % call: d["Capacity_Loss"],
                    "out_Reconstruction_Loss_%d" % call: d["Reconstruction_Loss"], "grad_recons_%d" % call: rl.grad,
                    "grad_mu_%d" % call: mul.grad, "grad_log_var_%d" % call: lvl.grad, "grad_q_%d" % call: ql.grad})
    assert m.num_iter == 3
    save("jointvae", **out)


def gen_last_models():
    """SURVEY 8(f) rank 2 / rank 4, last three: MSSIMVAE, DFCVAE and VampVAE loss functions of the unmodified reference."""
    g = torch.Generator().manual_seed(909)
    # MSSIMVAE: 32x32 so that the five dyadic levels (32 .. 2) and the trailing pool exist
    r, x = images(g, 4, hw=32)
    mu, lv = gaussians(g, 4, 16)
    m = ref.MSSIMVAE(3, 16)
    rl, mul, lvl = leaf(r), leaf(mu), leaf(lv)
    d = m.loss_function(rl, x, mul, lvl, M_N=0.005)
    d["loss"].backward()
    save("mssimvae", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005, out_loss=d["loss"],
         out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"], grad_recons=rl.grad, grad_mu=mul.grad,
         grad_log_var=lvl.grad)
    # DFCVAE: the constructor downloads VGG weights, and loss_function only reads alpha / beta -> bare instance
    m = ref.DFCVAE.__new__(ref.DFCVAE)
    torch.nn.Module.__init__(m)
    m.alpha, m.beta = 1.0, 0.5
    r, x = images(g, 5)
    mu, lv = gaussians(g, 5, 16)
    shapes = [(5, 8, 8, 8), (5, 12, 4, 4), (5, 16, 2, 2), (5, 16, 1, 1)]
    fr = [torch.relu(torch.randn(*sh, generator=g)) for sh in shapes]
    fi = [torch.relu(torch.randn(*sh, generator=g)) for sh in shapes]
    rl, mul, lvl, frl = leaf(r), leaf(mu), leaf(lv), [leaf(t) for t in fr]
    d = m.loss_function(rl, x, frl, fi, mul, lvl, M_N=0.005)
    d["loss"].backward()
    feats = {}
    for n, (a, b, al) in enumerate(zip(fr, fi, frl)):
        feats["in_fr%d" % n], feats["in_fi%d" % n], feats["grad_fr%d" % n] = a, b, al.grad
    save("dfcvae", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, hp_M_N=0.005, hp_alpha=1.0, hp_beta=0.5,
         out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         out_feature_loss=d["feature_loss"], grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad, **feats)
    # VampVAE: the pseudo-input encodings are injected (encode patched on the instance) so the fixture carries the
    # [K, D] prior and its gradients; Tensor.cuda is mapped to Tensor.to for this CPU run (vampvae.py:145).
    K, D, B = 10, 16, 7
    m = ref.VampVAE(3, D, num_components=K)
    r, x = images(g, B)
    mu, lv = gaussians(g, B, D)
    pmu, plv = gaussians(g, K, D)
    z = mu + torch.exp(0.5 * lv) * torch.randn(B, D, generator=g)
    rl, mul, lvl, zl, pml, pvl = leaf(r), leaf(mu), leaf(lv), leaf(z), leaf(pmu), leaf(plv)
    m.encode = lambda t: [pml, pvl]
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, device=None, **kw: self.to(device)
    try:
        d = m.loss_function(rl, x, mul, lvl, zl, M_N=0.005)
    finally:
        torch.Tensor.cuda = orig
    d["loss"].backward()
    save("vampvae", in_recons=r, in_input=x, in_mu=mu, in_log_var=lv, in_z=z, in_prior_mu=pmu, in_prior_log_var=plv,
         hp_M_N=0.005, out_loss=d["loss"], out_Reconstruction_Loss=d["Reconstruction_Loss"], out_KLD=d["KLD"],
         grad_recons=rl.grad, grad_mu=mul.grad, grad_log_var=lvl.grad, grad_z=zl.grad, grad_prior_mu=pml.grad,
         grad_prior_log_var=pvl.grad)


def gen_iw_forward():
    """IWAE.forward / MIWAE.forward / generate (iwae.py:121-127,188, miwae.py:124-130,192) of the reference classes in
    eval mode on a narrow conv stack; the module weights travel in the fixture (w_<state_dict key>).  IWAE.forward
    itself raises on current torch (iwae.py:103 view() of the permuted z), so its body is replayed line by
    line with the reference's own methods and a .contiguous() z handed to decode -- the same values."""
    g = torch.Generator().manual_seed(1290)
    B, D, K, M = 3, 8, 4, 2
    hd = [2, 4, 4, 8, 512]                                   # decode() hard-wires 512 channels (iwae.py:104)
    x = torch.rand(B, 3, 64, 64, generator=g)
    torch.manual_seed(1290)
    m = ref.IWAE(3, D, hidden_dims=list(hd), num_samples=K).eval()
    eps = torch.randn(B, K, D, generator=g)
    with torch.no_grad():
        mu, lv = m.encode(x)
        mu = mu.repeat(K, 1, 1).permute(1, 0, 2)
        lv = lv.repeat(K, 1, 1).permute(1, 0, 2)
        with injected_noise(eps):
            z = m.reparameterize(mu, lv)
        e = (z - mu) / lv
        rec = m.decode(z.contiguous())
    save("iwae_forward", in_input=x, noise_eps=eps, hp_K=K, hp_D=D, out_recons=rec, out_mu=mu.contiguous(),
         out_log_var=lv.contiguous(), out_z=z.contiguous(), out_eps=e.contiguous(), out_generate=rec[:, 0, :],
         **{"w_" + k: v for k, v in m.state_dict().items()})
    torch.manual_seed(1291)
    m = ref.MIWAE(3, D, hidden_dims=list(hd), num_samples=K, num_estimates=M).eval()
    eps = torch.randn(B, M, K, D, generator=g)
    with torch.no_grad():
        with injected_noise(eps):
            rec, xin, mu, lv, z, e = m(x)
        with injected_noise(eps):
            gen = m.generate(x)
    assert xin is x
    save("miwae_forward", in_input=x, noise_eps=eps, hp_K=K, hp_M=M, hp_D=D, out_recons=rec, out_mu=mu.contiguous(),
         out_log_var=lv.contiguous(), out_z=z.contiguous(), out_eps=e.contiguous(), out_generate=gen,
         **{"w_" + k: v for k, v in m.state_dict().items()})


if __name__ == "__main__":
    torch.set_num_threads(4)
    only = sys.argv[1:]
    for fn in (gen_vanilla, gen_beta, gen_logcosh, gen_iw, gen_wae, gen_betatc, gen_swae, gen_vq, gen_info, gen_next_models, gen_cat, gen_joint,
               gen_last_models, gen_iw_forward):
        if not only or fn.__name__[4:] in only:
            fn()
