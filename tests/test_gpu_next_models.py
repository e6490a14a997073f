"""SURVEY 8(f) rank 2 drop-ins (ConditionalVAE, DIPVAE, FactorVAE): loss_function on the K1 / K2 kernels against golden
fixtures generated from the unmodified reference (cvae.py:131-146, dip_vae.py:124-164, fvae.py:148-201), plus the
forward / loss-dict contract of each class.  Tolerance 1e-5 relative."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
from goldens import assert_close, leaf, load  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-5
DEV = "cuda"


@pytest.fixture(scope="module")
def models():
    import vaekit
    from vaekit.models import vae_models
    vaekit.lib()
    return vae_models


def test_cvae(models):
    g = load("cvae", DEV)
    m = models["ConditionalVAE"](3, 10, 16).to(DEV)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert d["KLD"].requires_grad and d["Reconstruction_Loss"].requires_grad          # not detached, unlike VanillaVAE
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)
    # forward contract: labels are required, four outputs, decoder sees [z | y]
    x = torch.rand(4, 3, 64, 64, device=DEV)
    y = torch.nn.functional.one_hot(torch.tensor([1, 3, 5, 7]), 10).to(DEV)
    out = m(x, labels=y)
    assert len(out) == 4 and out[0].shape == x.shape and out[2].shape == (4, 16)
    m.loss_function(*out, M_N=0.005)["loss"].backward()
    assert m.sample(4, 0, labels=y).shape == x.shape
    with pytest.raises(KeyError):
        m(x)


def test_dipvae(models):
    g = load("dipvae", DEV)
    m = models["DIPVAE"](3, 16, lambda_diag=10., lambda_offdiag=5.).to(DEV)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "DIP_Loss"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)
    out = m(torch.rand(4, 3, 64, 64, device=DEV))
    assert len(out) == 4
    m.loss_function(*out, M_N=0.005)["loss"].backward()


def test_factorvae(models):
    g = load("factorvae", DEV)
    m = models["FactorVAE"](3, 16, gamma=40.).to(DEV)
    lin = torch.nn.Linear(16, 2, bias=False).to(DEV)
    with torch.no_grad():
        lin.weight.copy_(g["in_W"].t())
    m.discriminator = lin                                     # the fixture's stand-in discriminator
    r, mu, lv, z = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_z"])
    d = m.loss_function(r, g["in_input"], mu, lv, z, M_N=g["hp_M_N"], optimizer_idx=0)
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "VAE_TC_Loss"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(z.grad, g["grad_z"], TOL); assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)
    # both optimizer branches through the full model (fvae.py:165-197)
    m2 = models["FactorVAE"](3, 16).to(DEV)
    out = m2(torch.rand(4, 3, 64, 64, device=DEV))
    assert len(out) == 5
    d0 = m2.loss_function(*out, M_N=0.005, optimizer_idx=0)
    assert set(d0) == {"loss", "Reconstruction_Loss", "KLD", "VAE_TC_Loss"}
    d0["loss"].backward(retain_graph=True)
    d1 = m2.loss_function(*out, M_N=0.005, optimizer_idx=1)
    assert set(d1) == {"loss", "D_TC_Loss"}
    d1["loss"].backward()


@pytest.mark.parametrize("shape", [(6, 3, 16, 16), (5, 3, 7, 9), (300, 3, 64, 64), (1, 1, 1, 1)])
def test_per_image_mse_and_step_report(models, shape):
    """the trainer's per-image MSE (experiment.py:58-62) as one streaming kernel and the whole step
    read-back (loss dict + per-image values) as one device-to-host copy."""
    from vaekit import functional as F
    gen = torch.Generator().manual_seed(sum(shape))
    r = torch.tanh(torch.randn(*shape, generator=gen)).to(DEV); x = torch.rand(*shape, generator=gen).to(DEV)
    got = F.per_image_mse(r, x)
    ref = torch.nn.functional.mse_loss(r.double(), x.double(), reduction="none").mean(dim=[1, 2, 3])
    assert got.shape == (shape[0],)
    assert_close(got, ref, 2e-6, "per-image mse")
    d = {"loss": torch.tensor([1.5], device=DEV), "KLD": torch.tensor(-2.25, device=DEV)}       # Beta-B style [1] loss
    vals, per = F.step_report(d, got)
    assert vals == {"loss": 1.5, "KLD": -2.25} and len(per) == shape[0]
    assert abs(per[0] - ref[0].item()) <= 2e-6 * abs(ref[0].item()) + 1e-12
    vals, per = F.step_report(d)
    assert per is None and vals["KLD"] == -2.25


def test_categorical_vae_golden(models):
    """CategoricalVAE on vk_gumbel_softmax_* / vk_cat_kl_rows_* against the reference's fixture
    (uniforms injected), forward and backward."""
    from vaekit import functional as F
    g = load("catvae", DEV)
    q = leaf(g["in_q"])
    s, _ = F.gumbel_softmax(q, g["hp_temp"], 1e-7, uniforms=g["noise_u"])
    s.view(5, 240).backward(g["in_gs"])
    assert_close(s.view(5, 240), g["out_s"], TOL, "gumbel-softmax sample")
    assert_close(q.grad, g["grad_s_q"], TOL, "d sample / d logits")
    m = models["CategoricalVAE"](3, 6, categorical_dim=40, temperature=0.5, alpha=30.).to(DEV).eval()
    r, q = leaf(g["in_recons"]), leaf(g["in_q"])
    d = m.loss_function(r, g["in_input"], q, M_N=g["hp_M_N"], batch_idx=3)
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(q.grad, g["grad_q"], TOL, "d loss / d logits")


@pytest.mark.parametrize("shape", [(3, 7, 40), (2, 5, 33), (4, 3, 1), (2, 2, 130), (64, 128, 40)])
def test_gumbel_softmax_own_noise_vs_oracle(models, shape):
    """The kernel's own Philox uniforms (emitted on request) fed to the oracle: same sample, same gradient; the draw is
    a pure function of (seed, offset) and lies strictly inside (0, 1)."""
    import vae_oracle as O
    from vaekit import functional as F
    gen = torch.Generator().manual_seed(sum(shape))
    q0 = 2.0 * torch.randn(*shape, generator=gen)
    gs = torch.randn(shape[0], shape[1] * shape[2], generator=gen)
    q = leaf(q0.to(DEV))
    s, u = F.gumbel_softmax(q, 0.7, 1e-7, rng=(1265, 40), want_uniforms=True)
    s2, u2 = F.gumbel_softmax(q0.to(DEV), 0.7, 1e-7, rng=(1265, 40), want_uniforms=True)
    assert torch.equal(u, u2) and torch.equal(s, s2)
    assert 0.0 < u.min().item() and u.max().item() < 1.0
    s.view(shape[0], -1).backward(gs.to(DEV))
    qc = leaf(q0)
    ref = O.gumbel_softmax(qc, u.cpu(), 0.7)
    ref.backward(gs)
    assert_close(s.view(shape[0], -1), ref, TOL, "sample")
    assert_close(q.grad, qc.grad, TOL, "gradient")
    rows = F.cat_kl_rows(leaf(q0.to(DEV)))
    p = torch.softmax(q0.double(), -1)
    import math
    want = (p * torch.log(p + 1e-7) - p * math.log(1. / shape[2] + 1e-7)).sum(dim=(1, 2))
    assert torch.allclose(rows.double().cpu(), want, rtol=1e-5, atol=1e-6 * shape[1])


def test_categorical_vae_contract(models):
    """forward list, annealing of the temperature every anneal_interval batches in training mode only, sample()."""
    m = models["CategoricalVAE"](3, 8, categorical_dim=10, anneal_interval=2, anneal_rate=0.5, temperature=0.5).to(DEV)
    x = torch.rand(4, 3, 64, 64, device=DEV)
    out = m(x)
    assert len(out) == 3 and out[0].shape == x.shape and out[2].shape == (4, 8, 10)
    m.temp = 2.0
    m.train()
    m.loss_function(*out, M_N=0.005, batch_idx=2)["loss"].backward()
    assert abs(float(m.temp) - max(2.0 * 2.718281828459045 ** (-1.0), 0.5)) < 1e-6      # annealed: batch_idx % 2 == 0
    t = float(m.temp)
    m.loss_function(*m(x), M_N=0.005, batch_idx=3)
    assert float(m.temp) == t                                                           # not on odd batches
    m.eval()
    m.loss_function(*m(x), M_N=0.005, batch_idx=4)
    assert float(m.temp) == t                                                           # not in eval mode
    assert m.sample(3, 0).shape == (3, 3, 64, 64)


def test_joint_vae_golden(models):
    """JointVAE on K1 + K8 against the reference's fixture: the Gumbel half of reparameterize with the
    fixture's uniforms, the Gaussian half through the noise the kernel itself reports, and two training-mode loss calls
    (the capacity targets move with num_iter)."""
    from vaekit import functional as F
    import vae_oracle as O
    g = load("jointvae", DEV)
    q = leaf(g["in_q"])
    s, _ = F.gumbel_softmax(q, 0.5, 1e-7, uniforms=g["noise_u"])
    s.backward(g["in_gz"][:, 12:])
    assert_close(s, g["out_z"][:, 12:], TOL, "categorical half of z")
    assert_close(q.grad, g["grad_z_q"], TOL, "d z / d q")
    m = models["JointVAE"](3, 12, 10, latent_num_iter=50, categorical_num_iter=20).to(DEV).train()
    z = m.reparameterize(leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_q"]))
    assert z.shape == (6, 22) and torch.allclose(z[:, 12:].sum(1), torch.ones(6, device=DEV), atol=1e-5)
    for call in (1, 2):
        r, mu, lv, q = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_q"])
        assert m.num_iter == call
        d = m.loss_function(r, g["in_input"], q, mu, lv, M_N=g["hp_M_N"], batch_idx=7)
        d["loss"].backward()
        for k in ("loss", "Capacity_Loss", "Reconstruction_Loss"):
            assert_close(d[k], g["out_%s_%d" % (k, call)], TOL, "%s call %d" % (k, call))
        assert_close(q.grad, g["grad_q_%d" % call], TOL); assert_close(mu.grad, g["grad_mu_%d" % call], TOL)
        assert_close(lv.grad, g["grad_log_var_%d" % call], TOL); assert_close(r.grad, g["grad_recons_%d" % call], TOL)
    m.eval()
    m.loss_function(leaf(g["in_recons"]), g["in_input"], leaf(g["in_q"]), leaf(g["in_mu"]), leaf(g["in_log_var"]),
                    M_N=0.005, batch_idx=7)
    assert m.num_iter == 3                                      # eval mode leaves the counter alone
    out = m(torch.rand(4, 3, 64, 64, device=DEV))
    assert len(out) == 5 and out[2].shape == (4, 10)
    assert m.sample(3, 0).shape == (3, 3, 64, 64)


def test_twostage_vae(models):
    g = load("twostage", DEV)
    m = models["TwoStageVAE"](3, 16).to(DEV)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert d["KLD"].requires_grad
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)
    out = m(torch.rand(4, 3, 64, 64, device=DEV))
    m.loss_function(*out, M_N=0.005)["loss"].backward()
    assert m.encoder2[0][0].weight.grad is None                 # the second stage is never called, as in the reference


def test_mssim_vae(models):
    """MSSIMVAE.loss_function (K1 rows + the torch MS-SSIM pyramid) against the reference fixture."""
    g = load("mssimvae", DEV)
    m = models["MSSIMVAE"](3, 16).to(DEV)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False                          # the fixture is an fp32 CPU run of the reference
    try:
        d = m.loss_function(r, g["in_input"], mu, lv, M_N=g["hp_M_N"])
        d["loss"].backward()
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    assert_close(d["KLD"], g["out_KLD"], TOL, "KLD")
    for k in ("loss", "Reconstruction_Loss"):                        # torch / cuDNN depthwise convs, not a vaekit kernel
        assert_close(d[k], g["out_" + k], 5e-5, k)
    assert d["KLD"].requires_grad
    assert_close(r.grad, g["grad_recons"], 5e-5, "recons")
    assert_close(mu.grad, g["grad_mu"], TOL); assert_close(lv.grad, g["grad_log_var"], TOL)
    out = m(torch.rand(4, 3, 64, 64, device=DEV))
    assert len(out) == 4
    m.loss_function(*out, M_N=0.005)["loss"].backward()
    assert m.fc_mu.weight.grad is not None and m.sample(3, 0).shape == (3, 3, 64, 64)


def test_dfc_vae(models):
    """DFCVAE.loss_function (pixel + four feature-map K2 reductions, K1 rows) against the reference
    fixture, and the forward contract with a stand-in extractor (the VGG weights are a download in the reference)."""
    g = load("dfcvae", DEV)
    feats = torch.nn.Sequential(*[torch.nn.Conv2d(3, 3, 3, padding=1) if i % 3 == 0 else torch.nn.ReLU() for i in range(44)])
    net = torch.nn.Module()
    net.features = feats
    m = models["DFCVAE"](3, 16, alpha=g["hp_alpha"], beta=g["hp_beta"], feature_network=net).to(DEV)
    r, mu, lv = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"])
    fr = [leaf(g["in_fr%d" % n]) for n in range(4)]
    fi = [g["in_fi%d" % n] for n in range(4)]
    d = m.loss_function(r, g["in_input"], fr, fi, mu, lv, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD", "feature_loss"):
        assert_close(d[k], g["out_" + k], TOL, k)
    assert_close(r.grad, g["grad_recons"], TOL); assert_close(mu.grad, g["grad_mu"], TOL)
    assert_close(lv.grad, g["grad_log_var"], TOL)
    for n in range(4):
        assert_close(fr[n].grad, g["grad_fr%d" % n], TOL, "feature %d" % n)
    out = m(torch.rand(4, 3, 64, 64, device=DEV))
    assert len(out) == 6 and len(out[2]) == 4 and len(out[3]) == 4
    m.loss_function(*out, M_N=0.005)["loss"].backward()
    assert m.fc_mu.weight.grad is not None and all(p.grad is None for p in m.feature_network.parameters())


def test_vamp_vae_golden(models):
    """VampVAE.loss_function on vk_gauss_logq_rows_* / vk_vamp_prior_* / K2 against the reference
    fixture (pseudo-input encodings injected), forward and backward into all six inputs."""
    g = load("vampvae", DEV)
    m = models["VampVAE"](3, 16, num_components=10).to(DEV)
    r, mu, lv, z = leaf(g["in_recons"]), leaf(g["in_mu"]), leaf(g["in_log_var"]), leaf(g["in_z"])
    pm, pv = leaf(g["in_prior_mu"]), leaf(g["in_prior_log_var"])
    m.prior_components = lambda inp: [pm, pv]
    d = m.loss_function(r, g["in_input"], mu, lv, z, M_N=g["hp_M_N"])
    d["loss"].backward()
    for k in ("loss", "Reconstruction_Loss", "KLD"):
        assert_close(d[k], g["out_" + k], TOL, k)
    for t, n in ((r, "recons"), (mu, "mu"), (lv, "log_var"), (z, "z"), (pm, "prior_mu"), (pv, "prior_log_var")):
        assert_close(t.grad, g["grad_" + n], TOL, n)
    # full model: gradients reach the pseudo-input embedding through the encoder
    m2 = models["VampVAE"](3, 16, num_components=10).to(DEV)
    out = m2(torch.rand(4, 3, 64, 64, device=DEV))
    assert len(out) == 5 and out[4].shape == (4, 16)
    d = m2.loss_function(*out, M_N=0.005)
    assert set(d) == {"loss", "Reconstruction_Loss", "KLD"}
    d["loss"].backward()
    assert m2.embed_pseudo[0].weight.grad is not None and torch.isfinite(m2.embed_pseudo[0].weight.grad).all()


@pytest.mark.parametrize("B,K,D", [(7, 10, 16), (144, 50, 128), (64, 500, 128), (5, 1, 3), (33, 37, 70), (300, 3, 300)])
def test_vamp_prior_kernels_vs_oracle(models, B, K, D):
    """K9 kernels against the float64 oracle (vampvae.py:136-158) on shapes around the shipped config (B 64 / 144, K 50 /
    500, D 128) and ragged ones; values and all gradients, plus run-to-run bit reproducibility (no atomics)."""
    import vae_oracle as O
    from vaekit import functional as F
    gen = torch.Generator().manual_seed(B * 1000 + K * 10 + D)
    z, mu = torch.randn(B, D, generator=gen), torch.randn(B, D, generator=gen)
    lv = 0.5 * torch.randn(B, D, generator=gen) - 1.0
    pm, pv = torch.randn(K, D, generator=gen), 0.5 * torch.randn(K, D, generator=gen) - 0.5
    gq, gp = torch.randn(B, generator=gen), torch.randn(B, generator=gen)
    dev = [leaf(t.to(DEV)) for t in (z, mu, lv, pm, pv)]
    ref = [leaf(t.double()) for t in (z, mu, lv, pm, pv)]
    lq = F.gauss_logq_rows(dev[0], dev[1], dev[2])
    lp = F.vamp_prior_rows(dev[0], dev[3], dev[4])
    (lq * gq.to(DEV)).sum().backward(retain_graph=True)
    gz_q = dev[0].grad.clone(); dev[0].grad = None
    (lp * gp.to(DEV)).sum().backward(retain_graph=True)
    lq_o = O.vamp_logq_rows(ref[0], ref[1], ref[2])
    lp_o = O.vamp_prior_rows(ref[0], ref[3], ref[4])
    (lq_o * gq.double()).sum().backward()
    gz_q_o = ref[0].grad.clone(); ref[0].grad = None
    (lp_o * gp.double()).sum().backward()
    tol = 2e-5
    assert_close(lq, lq_o, tol, "log q rows"); assert_close(lp, lp_o, tol, "vamp prior rows")
    assert_close(gz_q, gz_q_o, tol, "d logq / d z")
    for a, b, n in zip(dev, ref, ("z", "mu", "log_var", "prior_mu", "prior_log_var")):
        assert_close(a.grad, b.grad, tol, "grad " + n)
    first = [t.grad.clone() for t in dev]
    for t in dev:
        t.grad = None
    (lp * gp.to(DEV)).sum().backward()
    assert torch.equal(dev[0].grad, first[0]) and torch.equal(dev[3].grad, first[3]) and torch.equal(dev[4].grad, first[4])
    with pytest.raises(RuntimeError):
        F.vamp_prior_rows(dev[0], dev[3][:, :-1], dev[4][:, :-1])
