"""Parity of the stand-alone kernel-matrix methods (vk_pair_kernel_*: WAE_MMD / InfoVAE compute_kernel, compute_rbf,
compute_inv_mult_quad called directly; reference wae_mmd.py:125-191, info_vae.py:150-201) against the CPU oracle,
which tests/test_oracle_pin.py pins to the live reference.  Values and gradients within 1e-5 relative (fp32)."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import vae_oracle as O  # noqa: E402
from goldens import assert_close, leaf  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-5
DEV = "cuda"


@pytest.fixture(scope="module")
def vk():
    import vaekit
    from vaekit.models import vae_models
    vaekit.lib()
    return vae_models


@pytest.mark.parametrize("cls", ["WAE_MMD", "InfoVAE"])
@pytest.mark.parametrize("kt", ["rbf", "imq"])
@pytest.mark.parametrize("N,D", [(33, 24), (128, 128), (7, 3), (200, 130)])
def test_compute_kernel(vk, cls, kt, N, D):
    g = torch.Generator().manual_seed(N + D)
    m = vk[cls](3, D, kernel_type=kt, latent_var=1.5)
    x1, x2 = torch.randn(N, D, generator=g) + 0.2, 0.8 * torch.randn(N, D, generator=g)
    w = torch.randn(N, N, generator=g)                       # upstream weights for the rbf matrix
    a1, a2 = leaf(x1.to(DEV)), leaf(x2.to(DEV))
    out = m.compute_kernel(a1, a2)
    b1, b2 = leaf(x1), leaf(x2)
    ref = O.rbf_kernel_matrix(b1, b2, 1.5) if kt == "rbf" else O.imq_kernel_offdiag_sum(b1, b2, 1.5)
    assert out.shape == ref.shape
    assert_close(out, ref, TOL, "kernel value")
    if kt == "rbf":
        (out * w.to(DEV)).sum().backward(); (ref * w).sum().backward()
    else:
        (out * 1.3).backward(); (ref * 1.3).backward()
    assert_close(a1.grad, b1.grad, TOL, "d/dx1")
    assert_close(a2.grad, b2.grad, TOL, "d/dx2")


@pytest.mark.parametrize("kt", ["rbf", "imq"])
def test_same_tensor_both_sides(vk, kt):
    """compute_kernel(z, z) as compute_mmd calls it (wae_mmd.py:198): both roles' gradients add up on one leaf."""
    g = torch.Generator().manual_seed(3)
    m = vk["WAE_MMD"](3, 16, kernel_type=kt, latent_var=2.0)
    z = torch.randn(40, 16, generator=g)
    a, b = leaf(z.to(DEV)), leaf(z)
    out = m.compute_kernel(a, a)
    ref = O.rbf_kernel_matrix(b, b, 2.0) if kt == "rbf" else O.imq_kernel_offdiag_sum(b, b, 2.0)
    out.mean().backward(); ref.mean().backward()
    assert_close(out, ref, TOL, "value")
    assert_close(a.grad, b.grad, TOL, "grad")


@pytest.mark.parametrize("kt", ["rbf", "imq"])
def test_dense_operands(vk, kt):
    """compute_rbf / compute_inv_mult_quad on general [N1, N2, D] tensors (no broadcast structure), one of them a
    non-contiguous view."""
    g = torch.Generator().manual_seed(4)
    m = vk["WAE_MMD"](3, 20, kernel_type=kt, latent_var=0.7)
    d1 = torch.randn(6, 9, 20, generator=g); d2 = torch.randn(9, 6, 20, generator=g)
    a1, a2 = leaf(d1.to(DEV)), leaf(d2.to(DEV))
    b1, b2 = leaf(d1), leaf(d2)
    if kt == "rbf":
        out, ref = m.compute_rbf(a1, a2.transpose(0, 1)), O.compute_rbf(b1, b2.transpose(0, 1), 0.7)
    else:
        out = m.compute_inv_mult_quad(a1, a2.transpose(0, 1), eps=1e-3)
        ref = O.compute_inv_mult_quad(b1, b2.transpose(0, 1), 0.7, eps=1e-3)
    assert out.shape == ref.shape
    assert_close(out, ref, TOL, "value")
    (out.sum() * 0.5).backward(); (ref.sum() * 0.5).backward()
    assert_close(a1.grad, b1.grad, TOL, "d/dx1")
    assert_close(a2.grad, b2.grad, TOL, "d/dx2")


def test_undefined_kernel_type(vk):
    m = vk["WAE_MMD"](3, 8, kernel_type="laplace")
    x = torch.randn(4, 8, device=DEV)
    with pytest.raises(ValueError, match="Undefined kernel type."):
        m.compute_kernel(x, x)


def test_mmd_from_kernel_matrices_equals_sweep(vk):
    """wae_mmd.py:193-204 assembled from the three stand-alone kernel matrices equals the fused tcgen05 sweep."""
    from vaekit import functional as F
    g = torch.Generator().manual_seed(6)
    z = (torch.randn(256, 32, generator=g) + 0.25).to(DEV); pz = torch.randn(256, 32, generator=g).to(DEV)
    for kt in ("rbf", "imq"):
        m = vk["WAE_MMD"](3, 32, kernel_type=kt, latent_var=2.0)
        reg = 0.01
        mmd = reg * m.compute_kernel(pz, pz).mean() + reg * m.compute_kernel(z, z).mean() - \
            2 * reg * m.compute_kernel(pz, z).mean()
        fused, terms = F.mmd(z, pz, reg, kt, 2.0)
        scale = reg * terms.abs().max().item()
        assert abs(mmd.item() - fused.item()) <= 1e-5 * scale, kt
