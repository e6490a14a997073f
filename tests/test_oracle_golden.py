"""The oracle (oracle/vae_oracle.py) against the committed golden fixtures
(tests/golden/*.npz, produced from the unmodified reference by
tests/golden/make_golden.py).  CPU-only; runs on any box."""
import sys, os

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import vae_oracle as O  # noqa: E402
from goldens import assert_close, leaf, load  # noqa: E402

TOL = 2e-6  # same torch ops on the same inputs: only reduction-order noise is allowed


def test_vanilla():
    g = load("vanilla")
    mu, lv = leaf(g["in_mu"]), leaf(g["in_log_var"])
    z = O.reparameterize(mu, lv, g["noise_eps"])
    z.backward(g["in_gz"])
    assert_close(z, g["out_z"], TOL, "z")
    assert_close(mu.grad, g["grad_z_mu"], TOL)
    assert_close(lv.grad, g["grad_z_log_var"], TOL)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = O.vanilla_loss(r, g["in_input"], mu, lv, g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert not d["KLD"].requires_grad and not d["Reconstruction_Loss"].requires_grad
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)


@pytest.mark.parametrize("lt", ["H", "B"])
def test_beta(lt):
    g = load("beta_" + lt)
    kw = dict(loss_type=lt, beta=10 if lt == "H" else 4, gamma=10.0, C_max=25, C_stop_iter=10000)
    for call in (1, 2, 3):
        r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
        d = O.beta_loss(r, g["in_input"], mu, lv, g["hp_M_N"], num_iter=call, **kw)
        d["loss"].sum().backward()
        assert d["loss"].shape == g[f"out_loss_{call}"].shape
        assert_close(d["loss"], g[f"out_loss_{call}"], TOL)
        assert_close(r.grad, g[f"grad_recons_{call}"], TOL)
        assert_close(mu.grad, g[f"grad_mu_{call}"], TOL)
        assert_close(lv.grad, g[f"grad_log_var_{call}"], TOL)
    assert_close(d["KLD"], g["out_KLD"], TOL)
    assert d["KLD"].requires_grad


def test_beta_B_sign_flip():
    g = load("beta_B_sign")
    for call in range(1, 6):
        r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
        d = O.beta_loss(r, g["in_input"], mu, lv, g["hp_M_N"], loss_type="B", beta=4, gamma=10.0,
                        C_max=25, C_stop_iter=4, num_iter=call)
        d["loss"].sum().backward()
        assert_close(d["loss"], g[f"out_loss_{call}"], TOL)
        assert_close(mu.grad, g[f"grad_mu_{call}"], TOL)


def test_beta_undefined_type():
    g = load("beta_H")
    with pytest.raises(ValueError, match="Undefined loss type"):
        O.beta_loss(g["in_recons"], g["in_input"], g["in_mu"], g["in_log_var"], 0.005, loss_type="X",
                    beta=1, gamma=1, C_max=25, C_stop_iter=1, num_iter=1)


def test_logcosh():
    g = load("logcosh")
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = O.logcosh_loss(r, g["in_input"], mu, lv, g["hp_M_N"], alpha=g["hp_alpha"], beta=g["hp_beta"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], TOL)


def _expand(g, name):
    if name == "miwae":
        f = lambda t: t.repeat(3, 5, 1, 1).permute(2, 0, 1, 3)
    else:
        f = lambda t: t.repeat(5, 1, 1).permute(1, 0, 2)
    return f(g["in_mu0"]), f(g["in_log_var0"])


@pytest.mark.parametrize("name", ["iwae", "iwae_w", "miwae"])
def test_iw(name):
    g = load(name)
    mu, lv = _expand(g, name)
    if "noise_eps" in g:
        assert_close(O.reparameterize(mu, lv, g["noise_eps"]), g["out_z"], TOL)
    r, mu, lv = leaf(g["in_recons"]), leaf(mu.contiguous()), leaf(lv.contiguous())
    d = O.iw_bound_loss(r, g["in_input"], mu, lv, g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)


@pytest.mark.parametrize("name,kt", [("wae_rbf_d16", "rbf"), ("wae_imq_d16", "imq"),
                                     ("wae_rbf_d128", "rbf"), ("wae_imq_d128", "imq")])
def test_wae(name, kt):
    g = load(name)
    r, z = leaf(g["in_recons"]), leaf(g["in_z"])
    d = O.wae_loss(r, g["in_input"], z, g["noise_prior"], reg_weight=g["hp_reg_weight"], kernel_type=kt,
                   z_var=g["hp_z_var"])
    d["loss"].backward()
    pp, zz, pz = O.mmd_terms(g["in_z"], g["noise_prior"], kt, g["hp_z_var"])
    assert_close(pp, g["out_term_pp"], TOL)
    assert_close(zz, g["out_term_zz"], TOL)
    assert_close(pz, g["out_term_pz"], TOL)
    # the MMD scalar is a cancellation: tolerance relative to the largest term
    n = g["in_z"].shape[0]
    reg = g["hp_reg_weight"] / (n * (n - 1))
    scale = reg * max(abs(pp.item()), abs(zz.item()), abs(pz.item()))
    assert abs(d["MMD"].item() - g["out_MMD"].item()) <= 1e-5 * scale
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(z.grad, g["grad_z"], 2e-5)
    # chunked restatement agrees with the literal form
    c = O.mmd_terms_chunked(g["in_z"], g["noise_prior"], kt, g["hp_z_var"], block=17)
    for a, b in zip(c, (pp, zz, pz)):
        assert_close(a, b, 5e-6)


def test_wae_undefined_kernel():
    g = load("wae_rbf_d16")
    with pytest.raises(ValueError, match="Undefined kernel type"):
        O.mmd(g["in_z"], g["noise_prior"], 1.0, "poly", 2.0)


@pytest.mark.parametrize("name", ["betatc_b16_d10", "betatc_b48_d32"])
def test_betatc(name):
    g = load(name)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    z = O.reparameterize(mu, lv, g["noise_eps"])
    assert_close(z, g["out_z"], TOL)
    d = O.betatc_loss(r, g["in_input"], mu, lv, z, g["hp_M_N"], alpha=1.0, beta=6.0, gamma=1.0,
                      anneal_rate=g["hp_anneal_train"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "TC_Loss", "MI_Loss"):
        assert_close(d[k], g["out_" + k], 5e-6, k)
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], 1e-5)
    assert_close(lv.grad, g["grad_log_var"], 1e-5)
    r, mu, lv, zl = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["out_z"])
    d = O.betatc_loss(r, g["in_input"], mu, lv, zl, g["hp_M_N"], alpha=1.0, beta=6.0, gamma=1.0, anneal_rate=1.0)
    d["loss"].backward()
    assert_close(d["loss"], g["out_loss_eval"], 5e-6)
    assert_close(zl.grad, g["grad_eval_z"], 1e-5)
    assert_close(mu.grad, g["grad_eval_mu"], 1e-5)
    assert_close(lv.grad, g["grad_eval_log_var"], 1e-5)


@pytest.mark.parametrize("name", ["swae_n64_d16", "swae_n256_d128"])
def test_swae(name):
    g = load(name)
    torch.manual_seed(int(g["hp_seed"]))
    proj = O.random_projections(g["in_z"].shape[1], g["in_proj"].shape[0])
    assert torch.equal(proj, g["in_proj"])          # CPU generator stream identical to the reference's
    r, z = leaf(g["in_recons"]), leaf(g["in_z"])
    d = O.swae_loss(r, g["in_input"], z, g["noise_prior"], proj, reg_weight=g["hp_reg_weight"], p=g["hp_p"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "SWD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(z.grad, g["grad_z"], TOL)


def test_swae_unknown_projection():
    with pytest.raises(ValueError, match="Unknown projection distribution"):
        O.random_projections(4, 4, "uniform")


@pytest.mark.parametrize("name", ["vq_default", "vq_normal"])
def test_vq(name):
    g = load(name)
    lat, cb = leaf(g["in_latents"]), leaf(g["in_codebook"])
    q, vq_loss, idx = O.vq_forward(lat, cb, g["hp_beta"])
    (vq_loss + (q * g["in_gq"]).sum()).backward()
    assert torch.equal(idx, g["out_indices"])
    assert_close(q, g["out_quantized"], TOL)
    assert_close(vq_loss, g["out_vq_loss"], TOL)
    assert_close(lat.grad, g["grad_latents"], TOL)
    assert_close(cb.grad, g["grad_codebook"], TOL)


@pytest.mark.parametrize("name,kt", [("info_imq_d16", "imq"), ("info_rbf_d128", "rbf")])
def test_infovae(name, kt):
    """info_vae.py:117-134 with z chained to (mu, log_var) through reparameterize."""
    g = load(name)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    z = O.reparameterize(mu, lv, g["noise_eps"])
    assert_close(z, g["out_z"], TOL)
    d = O.infovae_loss(r, g["in_input"], z, mu, lv, g["noise_prior"], g["hp_M_N"], alpha=g["hp_alpha"],
                       beta=g["hp_beta"], reg_weight=g["hp_reg_weight"], kernel_type=kt, z_var=g["hp_z_var"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    scale = max(abs(g["out_term_" + t].item()) for t in ("pp", "zz", "pz"))
    assert abs(d["MMD"].item() - g["out_MMD"].item()) <= 1e-5 * scale
    assert_close(r.grad, g["grad_recons"], TOL)
    assert_close(mu.grad, g["grad_mu"], 2e-5)
    assert_close(lv.grad, g["grad_log_var"], 2e-5)


def test_next_models():
    """ConditionalVAE, DIPVAE, FactorVAE (VAE branch)."""
    g = load("cvae")
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = O.cvae_loss(r, g["in_input"], mu, lv, g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert d["KLD"].requires_grad and d["Reconstruction_Loss"].requires_grad
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)
    g = load("dipvae")
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = O.dip_loss(r, g["in_input"], mu, lv, g["hp_M_N"], lambda_diag=10., lambda_offdiag=5.)
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "DIP_Loss"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)
    g = load("factorvae")
    r, mu, lv, z = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_z"])
    d = O.factor_vae_loss(r, g["in_input"], mu, lv, z @ g["in_W"], g["hp_M_N"], gamma=40.)
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "VAE_TC_Loss"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(z.grad, g["grad_z"], TOL); assert_close(mu.grad, g["grad_mu"], TOL)


def test_cat_vae():
    """Gumbel-softmax reparameterisation and categorical KL of cat_vae.py."""
    g = load("catvae")
    q = leaf(g["in_q"])
    s = O.gumbel_softmax(q, g["noise_u"], g["hp_temp"])
    s.backward(g["in_gs"])
    assert_close(s, g["out_s"], TOL, "s"); assert_close(q.grad, g["grad_s_q"], TOL, "ds/dq")
    r, q = leaf(g["in_recons"]), leaf(g["in_q"])
    d = O.cat_vae_loss(r, g["in_input"], q, g["hp_M_N"], alpha=g["hp_alpha"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(q.grad, g["grad_q"], TOL)


def test_joint_vae():
    """JointVAE reparameterize (Gaussian | Gumbel-softmax) and capacity loss, two training calls."""
    g = load("jointvae")
    mu, lv, q = leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_q"])
    z = torch.cat([O.reparameterize(mu, lv, g["noise_eps"]), O.gumbel_softmax(q.view(6, 1, 10), g["noise_u"].view(6, 1, 10), 0.5)], dim=1)
    z.backward(g["in_gz"])
    assert_close(z, g["out_z"], TOL); assert_close(mu.grad, g["grad_z_mu"], TOL); assert_close(q.grad, g["grad_z_q"], TOL)
    kw = dict(alpha=30., disc_min=0., disc_max=25., disc_iter=20, disc_gamma=30., cont_min=0., cont_max=25., cont_iter=50,
              cont_gamma=30.)
    for call in (1, 2):
        r, mu, lv, q = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_q"])
        d = O.joint_vae_loss(r, g["in_input"], q, mu, lv, g["hp_M_N"], num_iter=call, **kw)
        d["loss"].backward()
        for k in ("loss", "Capacity_Loss", "Reconstruction_Loss"):
            assert_close(d[k], g["out_%s_%d" % (k, call)], TOL, k)
        assert_close(q.grad, g["grad_q_%d" % call], TOL); assert_close(mu.grad, g["grad_mu_%d" % call], TOL)
        assert_close(lv.grad, g["grad_log_var_%d" % call], TOL); assert_close(r.grad, g["grad_recons_%d" % call], TOL)


def test_twostage_vae():
    g = load("twostage")
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = O.cvae_loss(r, g["in_input"], mu, lv, g["hp_M_N"])          # twostage_vae.py:141-167 is the same undetached ELBO
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)


def test_mssim_vae():
    g = load("mssimvae")
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = O.mssim_vae_loss(r, g["in_input"], mu, lv, g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)


def test_dfc_vae():
    g = load("dfcvae")
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    fr = [leaf(g["in_fr%d" % n]) for n in range(4)]
    fi = [g["in_fi%d" % n] for n in range(4)]
    d = O.dfc_loss(r, g["in_input"], fr, fi, mu, lv, g["hp_M_N"], alpha=g["hp_alpha"], beta=g["hp_beta"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "feature_loss"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)
    for n in range(4):
        assert_close(fr[n].grad, g["grad_fr%d" % n], TOL, "feature %d" % n)


def test_vamp_vae():
    g = load("vampvae")
    r, mu, lv, z = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_z"])
    pm, pv = leaf(g["in_prior_mu"]), leaf(g["in_prior_log_var"])
    d = O.vamp_loss(r, g["in_input"], mu, lv, z, pm, pv, g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    for t, n in ((r, "recons"), (mu, "mu"), (lv, "log_var"), (z, "z"), (pm, "prior_mu"), (pv, "prior_log_var")):
        assert_close(t.grad, g["grad_" + n], TOL, n)


def test_vamp_closed_form_backward():
    """The hand-derived gradients the K9 kernels implement (csrc/vamp.cu), restated in float64 numpy and checked against
    the reference's autograd gradients in the fixture: softmax weights from (t, lse), fixed-order sums."""
    g = load("vampvae", dtype=torch.float64)
    z, mu, lv = (g[k].double().numpy() for k in ("in_z", "in_mu", "in_log_var"))
    pm, pv = g["in_prior_mu"].double().numpy(), g["in_prior_log_var"].double().numpy()
    B, K, M_N = z.shape[0], pm.shape[0], g["hp_M_N"]
    diff = z[:, None, :] - pm[None]                                     # [B, K, D]
    e = np.exp(-pv)[None]
    t = np.sum(-0.5 * (pv[None] + diff ** 2) * e, axis=2) - np.log(K)
    lse = np.log(np.sum(np.exp(t - t.max(1, keepdims=True)), axis=1)) + t.max(1)
    w = np.exp(t - lse[:, None])
    g_lse = -M_N / B                                                    # d loss / d lse_i: kld = mean(logq) - mean(lse)
    g_lq = +M_N / B
    gz_prior = g_lse * np.einsum("bk,bkd->bd", w, -diff * e)
    g_pm = g_lse * np.einsum("bk,bkd->kd", w, diff * e)
    g_pv = g_lse * np.einsum("bk,bkd->kd", w, 0.5 * e * (pv[None] + diff ** 2 - 1.0))
    dq, eq = z - mu, np.exp(-lv)
    gz_q = g_lq * (-dq * eq)
    g_mu = g_lq * (dq * eq)
    g_lv = g_lq * 0.5 * eq * (lv + dq ** 2 - 1.0)
    assert_close(gz_prior + gz_q, g["grad_z"], TOL, "z")
    assert_close(g_pm, g["grad_prior_mu"], TOL, "prior_mu")
    assert_close(g_pv, g["grad_prior_log_var"], TOL, "prior_log_var")
    assert_close(g_mu, g["grad_mu"], TOL, "mu"); assert_close(g_lv, g["grad_log_var"], TOL, "log_var")
