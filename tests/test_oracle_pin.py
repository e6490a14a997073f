"""Pins oracle/vae_oracle.py against the LIVE reference imported read-only from
/root/reference (build container only; skipped on the GPU box, where the
committed fixtures of tests/test_oracle_golden.py take over)."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import vae_oracle as O  # noqa: E402
from goldens import assert_close, leaf  # noqa: E402
from refload import have_reference, injected_noise, load_reference_models  # noqa: E402

pytestmark = pytest.mark.skipif(not have_reference(), reason="/root/reference not present")
TOL = 2e-6


@pytest.fixture(scope="module")
def ref():
    return load_reference_models()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_full_vanilla_model_contract(ref, dtype):
    """tests/test_vae.py:11-28 of the reference: model(x) then loss_function(*result, M_N=0.005)."""
    torch.manual_seed(0)
    m = ref.VanillaVAE(3, 10).to(dtype)
    x = torch.randn(4, 3, 64, 64, dtype=dtype)
    eps = torch.randn(4, 10, dtype=dtype)
    with injected_noise(eps):
        res = m(x)
    assert [tuple(t.shape) for t in res] == [(4, 3, 64, 64), (4, 3, 64, 64), (4, 10), (4, 10)]
    d_ref = m.loss_function(*res, M_N=0.005)
    d = O.vanilla_loss(res[0], res[1], res[2], res[3], 0.005)
    for k in d_ref:
        assert_close(d[k], d_ref[k], TOL, k)


def test_logcosh_overflow_matches(ref):
    """logcosh_vae.py:146-150 overflows to inf for 2*alpha*|t| > 88.7; the oracle keeps that."""
    m = ref.LogCoshVAE(3, 8, alpha=100.0, beta=1.0)
    r = torch.full((2, 3, 4, 4), -1.0)
    x = torch.zeros(2, 3, 4, 4)
    mu = torch.zeros(2, 8); lv = torch.zeros(2, 8)
    a = m.loss_function(r, x, mu, lv, M_N=0.1)["loss"]
    b = O.logcosh_loss(r, x, mu, lv, 0.1, alpha=100.0, beta=1.0)["loss"]
    assert torch.isinf(a) and torch.isinf(b)


@pytest.mark.parametrize("kt", ["rbf", "imq"])
def test_mmd_random(ref, kt):
    g = torch.Generator().manual_seed(5)
    m = ref.WAE_MMD(3, 24, reg_weight=100, kernel_type=kt, latent_var=1.5)
    z = leaf(torch.randn(33, 24, generator=g)); pz = torch.randn(33, 24, generator=g)
    with injected_noise(pz):
        a = m.compute_mmd(z, 0.01)
    ga, = torch.autograd.grad(a, z)
    z2 = leaf(z.detach())
    b = O.mmd(z2, pz, 0.01, kt, 1.5)
    gb, = torch.autograd.grad(b, z2)
    assert abs(a.item() - b.item()) < 1e-7
    assert_close(gb, ga, 1e-5)


@pytest.mark.parametrize("kt", ["rbf", "imq"])
def test_kernel_matrix_methods(ref, kt):
    """compute_kernel / compute_rbf / compute_inv_mult_quad called directly (wae_mmd.py:125-191, info_vae.py:150-201)."""
    g = torch.Generator().manual_seed(8)
    for cls in (ref.WAE_MMD, ref.InfoVAE):
        m = cls(3, 12, reg_weight=100, kernel_type=kt, latent_var=1.5)
        x1 = leaf(torch.randn(9, 12, generator=g)); x2 = leaf(torch.randn(9, 12, generator=g))
        a = m.compute_kernel(x1, x2)
        ga = torch.autograd.grad(a.sum(), (x1, x2))
        y1, y2 = leaf(x1.detach()), leaf(x2.detach())
        b = O.rbf_kernel_matrix(y1, y2, 1.5) if kt == "rbf" else O.imq_kernel_offdiag_sum(y1, y2, 1.5)
        gb = torch.autograd.grad(b.sum(), (y1, y2))
        assert a.shape == b.shape and torch.equal(a, b)
        assert_close(gb[0], ga[0], 1e-6); assert_close(gb[1], ga[1], 1e-6)
        # dense [N1, N2, D] operands straight into the two kernel methods
        d1 = leaf(torch.randn(5, 7, 12, generator=g)); d2 = leaf(torch.randn(5, 7, 12, generator=g))
        a = m.compute_rbf(d1, d2) if kt == "rbf" else m.compute_inv_mult_quad(d1, d2)
        b = O.compute_rbf(d1, d2, 1.5) if kt == "rbf" else O.compute_inv_mult_quad(d1, d2, 1.5)
        assert a.shape == b.shape and torch.equal(a, b)


def test_zero_division_at_batch_one(ref):
    """wae_mmd.py:114-116 / swae.py:117-119: B=1 raises ZeroDivisionError."""
    m = ref.WAE_MMD(3, 8)
    with pytest.raises(ZeroDivisionError):
        m.loss_function(torch.zeros(1, 3, 4, 4), torch.zeros(1, 3, 4, 4), torch.zeros(1, 8))
    with pytest.raises(ZeroDivisionError):
        O.wae_loss(torch.zeros(1, 3, 4, 4), torch.zeros(1, 3, 4, 4), torch.zeros(1, 8), torch.zeros(1, 8),
                   reg_weight=100, kernel_type="imq", z_var=2.0)


def test_vq_random(ref):
    g = torch.Generator().manual_seed(9)
    q = ref.VectorQuantizer(64, 16, 0.25)
    lat = leaf(torch.randn(3, 16, 5, 7, generator=g))
    out, loss = q(lat)
    (loss + out.sum()).backward()
    lat2, cb = leaf(lat.detach()), leaf(q.embedding.weight.detach())
    o2, l2, _ = O.vq_forward(lat2, cb, 0.25)
    (l2 + o2.sum()).backward()
    assert torch.equal(out, o2)
    assert_close(l2, loss, TOL)
    assert_close(lat2.grad, lat.grad, TOL)
    assert_close(cb.grad, q.embedding.weight.grad, TOL)


def test_betatc_random(ref):
    g = torch.Generator().manual_seed(11)
    m = ref.BetaTCVAE(3, 12, anneal_steps=50); m.eval()
    r = torch.randn(20, 3, 4, 4, generator=g); x = torch.randn(20, 3, 4, 4, generator=g)
    mu = torch.randn(20, 12, generator=g); lv = 0.3 * torch.randn(20, 12, generator=g)
    z = torch.randn(20, 12, generator=g)
    a = m.loss_function(r, x, mu, lv, z, M_N=0.001)
    b = O.betatc_loss(r, x, mu, lv, z, 0.001, alpha=1.0, beta=6.0, gamma=1.0, anneal_rate=1.0)
    for k in a:
        assert_close(b[k], a[k], 5e-6, k)
