"""fp32-grade tcgen05 GEMM (vk_gemm_nt: three-way bf16 split, six passes, A operand in tensor memory) against float64.
It carries the SWAE projection (swae.py:164-166) and its backward, whose outputs are sort keys: the bar is the
accuracy of an fp32 FFMA GEMM, not of a bf16 one."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0) if torch.cuda.is_available() else None


@pytest.mark.parametrize("N,J,K", [(1, 1, 1), (256, 50, 128), (256, 128, 50), (1000, 200, 128), (4100, 128, 200),
                                   (130, 16, 16), (129, 17, 33), (70000, 200, 128), (300, 240, 128), (300, 64, 256)])
def test_gemm_nt_matches_float64(N, J, K):
    from vaekit import functional as F
    g = torch.Generator().manual_seed(N + 7 * J + 13 * K)
    a = torch.randn(N, K, generator=g) * 1.5 + 0.1
    b = torch.randn(J, K, generator=g)
    b = b / b.norm(dim=1, keepdim=True)
    ref = a.double() @ b.double().t()
    got = F.gemm_nt(a.to(DEV), b.to(DEV)).cpu()
    fp32 = (a @ b.t())
    err = (got.double() - ref).abs().max().item()
    err32 = (fp32.double() - ref).abs().max().item()
    scale = (a.double().abs() @ b.double().abs().t()).max().item()
    print("N=%d J=%d K=%d: max abs err %.3e (torch fp32 matmul %.3e), |A||B| scale %.2f" % (N, J, K, err, err32, scale))
    assert err <= 4e-7 * scale


def test_gemm_nt_strided_operands():
    from vaekit import functional as F
    g = torch.Generator().manual_seed(5)
    at = torch.randn(200, 3000, generator=g).to(DEV)            # A given k-major (the backward's g_lat[s][n])
    b = torch.randn(128, 200, generator=g).to(DEV)
    got = F.gemm_nt(at.t(), b)
    ref = at.t().double() @ b.double().t()
    assert (got.double() - ref).abs().max().item() <= 4e-7 * (at.t().double().abs() @ b.double().abs().t()).max().item()
    bt = torch.randn(64, 128, generator=g).to(DEV)               # B given as [K, J]
    a = torch.randn(500, 64, generator=g).to(DEV)
    got = F.gemm_nt(a, bt.t())
    ref = a.double() @ bt.double()
    assert (got.double() - ref).abs().max().item() <= 4e-7 * (a.double().abs() @ bt.double().abs()).max().item()


def test_gemm_nt_rejects_large_shapes():
    from vaekit import functional as F
    a = torch.zeros(8, 300, device=DEV)
    b = torch.zeros(8, 300, device=DEV)
    with pytest.raises(RuntimeError, match="do not fit"):
        F.gemm_nt(a, b)
