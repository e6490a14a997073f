"""vk_iw_bound_fwd_static: the allocation-free variant of the IWAE / MIWAE bound forward (scratch in a file-scope
`static __device__` array of the library instead of a caller workspace).  It has no counterpart in the reference; it was
added for callers that cannot call cudaMalloc.

Called through the C ABI (ctypes) and checked against the existing entry point vk_iw_bound_fwd on the same inputs, on
two calls in a row (the array keeps the previous call's contents), at the tolerance of the other streaming-kernel tests.
This file is named so that pytest collects it LAST: a failure here must not hide
the tests before it.  It passes alone in a fresh process and after the rest of the suite."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "pytorch-vae_b200"),):
    if p not in sys.path:
        sys.path.insert(0, p)

pytestmark = pytest.mark.gpu
TOL = 1e-5
DEV = "cuda"
MAX_B = 16384          # VK_IW_STATIC_MAX_BATCH (include/vaekit.h)


@pytest.fixture(scope="module")
def L():
    import vaekit
    from vaekit import _lib
    vaekit.lib()
    return _lib


def _inputs(B, G, K, P, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    r = torch.tanh(torch.randn(B, G, K, P, device=DEV, generator=g))
    x = torch.rand(B, P, device=DEV, generator=g)
    kl = 20.0 * torch.rand(B * G * K, device=DEV, generator=g)
    return r, x, kl


def _existing(L, r, x, kl, B, G, K, P, M_N):
    lp = torch.empty(B * G * K, dtype=torch.float32, device=DEV)
    out = torch.empty(4, dtype=torch.float32, device=DEV)
    ws = L.workspace(r.device, 256 + 24 * B)
    L.call("iw_bound_fwd", L.ptr(r), L.ptr(x), L.ptr(kl), B, G, K, P, M_N, L.ptr(lp), L.ptr(out), L.ptr(ws), ws.numel(),
           L.stream_ptr())
    return lp, out


def _static(L, r, x, kl, B, G, K, P, M_N):
    lp = torch.full((B * G * K,), float("nan"), dtype=torch.float32, device=DEV)
    out = torch.full((4,), float("nan"), dtype=torch.float32, device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    L.call("iw_bound_fwd_static", L.ptr(r), L.ptr(x), L.ptr(kl), B, G, K, P, M_N, L.ptr(lp), L.ptr(out), L.stream_ptr())
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before      # nothing went through torch's allocator
    return lp, out


def _rel(a, b):
    a, b = a.double().reshape(-1), b.double().reshape(-1)
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def test_static_scratch_entry_matches_existing_on_two_calls_in_a_row(L):
    # two different problems back to back (IWAE K=5, then MIWAE M=3 K=5 with a smaller batch: the second call finds the
    # first call's partials in the array), each against the existing entry point on the same inputs
    for B, G, K, P, M_N, seed in ((96, 1, 5, 3 * 16 * 16, 0.00025, 1), (40, 3, 5, 3 * 8 * 8, 0.05, 2)):
        r, x, kl = _inputs(B, G, K, P, seed)
        lp_ref, out_ref = _existing(L, r, x, kl, B, G, K, P, M_N)
        lp, out = _static(L, r, x, kl, B, G, K, P, M_N)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all() and torch.isfinite(lp).all()
        assert _rel(lp, lp_ref) <= TOL, (B, G, K)
        for i, name in enumerate(("loss", "mean lp", "mean kl", "-mean kl")):
            assert abs(out[i].item() - out_ref[i].item()) <= TOL * abs(out_ref[i].item()), (name, B, G, K)
    # the same call twice in a row gives the same answer (the ticket header is cleared per call, the partials rewritten)
    r, x, kl = _inputs(64, 1, 5, 3 * 16 * 16, 3)
    a = _static(L, r, x, kl, 64, 1, 5, 3 * 16 * 16, 0.00025)
    b = _static(L, r, x, kl, 64, 1, 5, 3 * 16 * 16, 0.00025)
    torch.cuda.synchronize()
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_static_scratch_entry_limit_and_arguments(L):
    # the largest batch the array holds works; a small call right after it sees 16384 stale partials behind its own 4
    B, G, K, P = MAX_B, 1, 2, 16
    r, x, kl = _inputs(B, G, K, P, 4)
    lp_ref, out_ref = _existing(L, r, x, kl, B, G, K, P, 0.01)
    lp, out = _static(L, r, x, kl, B, G, K, P, 0.01)
    torch.cuda.synchronize()
    assert _rel(lp, lp_ref) <= TOL and _rel(out, out_ref) <= TOL
    r2, x2, kl2 = _inputs(4, 1, 2, 16, 5)
    lp_ref, out_ref = _existing(L, r2, x2, kl2, 4, 1, 2, 16, 0.01)
    lp, out = _static(L, r2, x2, kl2, 4, 1, 2, 16, 0.01)
    torch.cuda.synchronize()
    assert _rel(lp, lp_ref) <= TOL and _rel(out, out_ref) <= TOL
    # one past the maximum: refused before anything is enqueued, with a message that names the limit
    lp = torch.zeros(8, device=DEV)
    out = torch.zeros(4, device=DEV)
    with pytest.raises(RuntimeError, match="static scratch maximum of %d" % MAX_B):
        L.call("iw_bound_fwd_static", L.ptr(r), L.ptr(x), L.ptr(kl), MAX_B + 1, 1, 2, 16, 0.01, L.ptr(lp), L.ptr(out),
               L.stream_ptr())
    with pytest.raises(RuntimeError, match="null pointer"):
        L.call("iw_bound_fwd_static", None, L.ptr(x), L.ptr(kl), 4, 1, 2, 16, 0.01, L.ptr(lp), L.ptr(out), L.stream_ptr())
    torch.cuda.synchronize()
    assert out.abs().sum().item() == 0.0                                        # the refused calls wrote nothing
