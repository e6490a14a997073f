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
