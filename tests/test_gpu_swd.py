"""Parity of the SWAE sliced-Wasserstein path (K6: projection GEMM, segmented radix sort, loss, d/dz) against
the golden fixtures made from the unmodified reference (models/swae.py:110-179) and against the CPU oracle.
Sorted values: bit-exact against torch.sort of the same keys; permutations: bit-exact wherever keys are not
tied (torch.sort is unstable by default, ours is stable); loss / gradient: 1e-5 relative."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import vae_oracle as O  # noqa: E402
from goldens import assert_close, leaf, load, rel_err  # noqa: E402

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


@pytest.mark.parametrize("segs,n", [(1, 1), (3, 7), (5, 4096), (4, 4097), (7, 10000), (2, 262144 + 13), (200, 5000)])
def test_seg_sort_matches_torch_sort(vk, segs, n):
    F, _ = vk
    gen = torch.Generator(device=DEV).manual_seed(segs * 1000 + n)
    keys = torch.randn(segs, n, device=DEV, generator=gen) * 3
    keys[0, : min(n, 5)] = torch.tensor([0.0, -0.0, 1e-38, -1e-38, 3.5])[: min(n, 5)].to(DEV)     # signed zeros, denormals
    if n >= 64:
        keys[-1, 10:40] = keys[-1, 10]                                                         # ties
    out, perm = F.seg_sort(keys, want_perm=True)
    ref, ref_perm = torch.sort(keys, dim=1, stable=True)
    assert torch.equal(out.view(torch.int32) & 0x7FFFFFFF | 0, ref.view(torch.int32) & 0x7FFFFFFF | 0) or torch.equal(out, ref)
    assert (out[:, 1:] >= out[:, :-1]).all()                                                    # sortedness
    assert torch.equal(torch.gather(keys, 1, perm), out)                                        # perm is a valid gather
    assert torch.equal(torch.sort(perm, dim=1).values, torch.arange(n, device=DEV).expand(segs, n))  # a permutation
    # stable order == torch's stable sort, except between +0.0 and -0.0 (distinct radix keys, equal floats)
    mism = (perm != ref_perm)
    if mism.any():
        assert (torch.gather(keys, 1, perm)[mism] == 0).all()
    out2, none = F.seg_sort(keys)
    assert none is None and torch.equal(out2, out)


def rank_clear_rows(z, proj, rel_gap=2e-6):
    """Rows none of whose projections has a float64 neighbour gap below rel_gap: two correct fp32 GEMMs (different
    summation order) may rank such near-tied samples differently, which moves their gradient by the neighbouring
    prior gap.  Parity of the gradient is asserted on the other rows; the rest is reported."""
    z, proj = z.detach().double().cpu(), proj.detach().double().cpu()
    n, S = z.shape[0], proj.shape[0]
    srt, order = torch.sort(z @ proj.t(), dim=0)
    gaps = srt[1:] - srt[:-1]
    big = torch.full((1, S), float("inf"), dtype=torch.float64)
    near = torch.minimum(torch.cat([gaps, big]), torch.cat([big, gaps])) < rel_gap * srt.abs().clamp_min(1.0)
    ambiguous = torch.zeros(n, dtype=torch.bool)
    for j in range(S):
        ambiguous[order[near[:, j], j]] = True
    return ~ambiguous


@pytest.mark.parametrize("name", ["swae_n64_d16", "swae_n256_d128"])
def test_swae_golden(vk, name):
    F, _ = vk
    g = load(name, DEV)
    n = g["in_z"].shape[0]
    reg = g["hp_reg_weight"] / (n * (n - 1))
    r, z = leaf(g["in_recons"]), leaf(g["in_z"])
    _, recons_loss, _, _ = F.elbo(r, g["in_input"], recon_mode=F.RECON_MSE_L1_MEAN)
    s = F.swd(z, g["noise_prior"], g["in_proj"], g["hp_p"], reg)
    (recons_loss + s).backward()
    assert_close(recons_loss, g["out_Reconstruction_Loss"], TOL)
    assert_close(s, g["out_SWD"], TOL, "SWD")
    assert_close(recons_loss + s, g["out_loss"], TOL)
    assert_close(r.grad, g["grad_recons"], TOL)
    clear = rank_clear_rows(g["in_z"], g["in_proj"])
    g_ours, g_ref = z.grad.cpu(), g["grad_z"].cpu()
    den = g_ref.abs().max().item()
    e_clear = (g_ours[clear] - g_ref[clear]).abs().max().item() / den
    e_all = (g_ours - g_ref).abs().max().item() / den
    print("%s: grad rel err %.2e on %d clear rows, %.2e overall (%d rank-ambiguous rows)" %
          (name, e_clear, int(clear.sum()), e_all, int((~clear).sum())))
    assert e_clear <= TOL, "grad z on rank-unambiguous rows: %.2e" % e_clear
    assert int((~clear).sum()) <= max(2, n // 20) and e_all <= 5e-3


@pytest.mark.parametrize("n,dim,S,p", [(1000, 128, 200, 2.0), (4100, 128, 50, 2.0), (333, 10, 7, 2.0), (2, 16, 3, 2.0),
                                       (5000, 64, 33, 2.0)])
def test_swd_vs_oracle(vk, n, dim, S, p):
    F, _ = vk
    gen = torch.Generator().manual_seed(n + S)
    z = 1.5 * torch.randn(n, dim, generator=gen) + 0.1
    prior = torch.randn(n, dim, generator=gen)
    torch.manual_seed(1265)
    proj = O.random_projections(dim, S)
    reg = 100.0 / (n * (n - 1))
    zc = leaf(z)
    ref = O.swd(zc, prior, proj, p, reg)
    ref.backward()
    zg = leaf(z.to(DEV))
    s = F.swd(zg, prior.to(DEV), proj.to(DEV), p, reg)
    (s * 0.7).backward()
    assert_close(s, ref, TOL, "swd")
    # the gradient goes through the sort permutation: see rank_clear_rows
    clear = rank_clear_rows(z, proj)
    ambiguous = ~clear
    g_ours, g_ref = (zg.grad / 0.7).cpu(), zc.grad
    den = g_ref.abs().max().item()
    e_all = (g_ours - g_ref).abs().max().item() / den
    e_clear = (g_ours[clear] - g_ref[clear]).abs().max().item() / den if clear.any() else 0.0
    print("swd n=%d d=%d S=%d: %.6e vs %.6e; grad rel err %.2e on %d clear rows, %.2e overall (%d rank-ambiguous rows)" %
          (n, dim, S, s.item(), ref.item(), e_clear, int(clear.sum()), e_all, int(ambiguous.sum())))
    assert e_clear <= TOL
    assert e_all <= 5e-3


def test_swd_projection_shards_add_up(vk):
    """Sharding by projection (the multi-GPU decomposition, emulated on one GPU)."""
    F, _ = vk
    gen = torch.Generator().manual_seed(4)
    n, dim, S = 3000, 128, 40
    z = (1.5 * torch.randn(n, dim, generator=gen) + 0.1).to(DEV)
    prior = torch.randn(n, dim, generator=gen).to(DEV)
    proj = O.random_projections(dim, S).to(DEV)
    tot, grad = F.swd_sweep(z, prior, proj, 2.0)
    t2 = torch.zeros(1, dtype=torch.float64, device=DEV)
    g2 = torch.zeros_like(grad)
    for a, b in ((0, 13), (13, 27), (27, 40)):
        t, g = F.swd_sweep(z, prior, proj[a:b], 2.0)
        t2 += t
        g2 += g
    assert_close(t2, tot, 1e-12)
    assert_close(g2, grad, 2e-6)


def test_swae_model_contract(vk):
    _, models = vk
    torch.manual_seed(1265)
    m = models["SWAE"](3, 128, reg_weight=100, num_projections=200).to(DEV)
    x = torch.rand(16, 3, 64, 64, device=DEV)
    out = m(x)
    d = m.loss_function(*out)                          # no M_N, like tests/test_swae.py:26
    assert set(d) == {"loss", "Reconstruction_Loss", "SWD"}
    d["loss"].backward()
    assert m.fc_z.weight.grad is not None
    ref = O.swae_loss(out[0].detach().cpu(), x.cpu(), out[2].detach().cpu(), m.last_prior_z.cpu(), m.last_proj.cpu(),
                      reg_weight=100, p=2.0)
    for k in d:
        assert_close(d[k], ref[k], TOL, k)
    # the projection directions follow the CPU generator exactly like the reference
    torch.manual_seed(7)
    a = m.get_random_projections(128, 200)
    torch.manual_seed(7)
    assert torch.equal(a, O.random_projections(128, 200))
    m.proj_dist = "uniform"
    with pytest.raises(ValueError, match="Unknown projection distribution"):
        m.get_random_projections(128, 4)
    with pytest.raises(ZeroDivisionError):
        m.loss_function(torch.rand(1, 3, 8, 8, device=DEV), torch.rand(1, 3, 8, 8, device=DEV), torch.randn(1, 128, device=DEV))


def dev_rel(a, b):
    """max |a - b| / max |b| evaluated on the device (the full-size tensors are GBs: no host round trip)."""
    b = b.double()
    return ((a.double() - b).abs().max() / b.abs().max().clamp_min(1e-300)).item()


def test_swd_full_size_stages(vk):
    """BASELINE configs[4] at its full size (N = 1 048 576, D = 128, S = 200), stage by stage against independent
    implementations on the same device data: projections vs a float64 matmul; segmented radix sort vs torch.sort (values
    and stable permutation bit-exact); loss vs a float64 sum over the torch-sorted segments; d/dlat vs the closed form
    scattered through torch's permutation; the fused vk_swd_fwd vs the stage composition; the gradient vs float64; and
    two size-independent properties: swd(z, z) = 0 exactly, and invariance under a permutation of the batch rows."""
    F, _ = vk
    n, dim, S, p = 1 << 20, 128, 200, 2.0
    gen = torch.Generator(device=DEV).manual_seed(5)
    z = 1.5 * torch.randn(n, dim, device=DEV, generator=gen) + 0.1
    prior = torch.randn(n, dim, device=DEV, generator=gen)
    torch.manual_seed(1265)
    proj = O.random_projections(dim, S).to(DEV)

    # 1. projection GEMM (fp32-grade tcgen05, three-way bf16 split) vs float64
    lat, pri = F.swd_project(z, proj), F.swd_project(prior, proj)
    lat64 = (z.double() @ proj.double().t()).t()
    e = dev_rel(lat, lat64)
    print("full-size projections: rel err vs fp64 %.2e" % e)
    assert e <= 1e-5
    del lat64

    # 2. segmented sort vs torch.sort on the same keys
    s_lat, perm = F.seg_sort(lat, want_perm=True)
    t_lat, t_perm = torch.sort(lat, dim=1, stable=True)
    assert torch.equal(s_lat, t_lat), "sorted keys differ from torch.sort"
    assert torch.equal(perm, t_perm), "permutation differs from torch's stable sort"
    del s_lat, perm
    eq = t_lat[:, 1:] == t_lat[:, :-1]                           # exact key ties, needed by step 5
    tie_sorted = torch.zeros_like(t_lat, dtype=torch.bool)
    tie_sorted[:, 1:] |= eq
    tie_sorted[:, :-1] |= eq
    tied = torch.zeros_like(tie_sorted).scatter_(1, t_perm, tie_sorted)      # sample order
    del eq, tie_sorted
    s_pri, _ = F.seg_sort(pri)
    t_pri = torch.sort(pri, dim=1)[0]
    assert torch.equal(s_pri, t_pri)
    del s_pri

    # 3. loss and d/dlat of the sort stage
    w = t_lat - t_pri                                            # fp32, like swae.py:176-178
    total_ref = (w.double() ** 2).sum().item()
    g_ref = torch.empty_like(lat).scatter_(1, t_perm, 2.0 * w)   # d sum / d lat[s, perm[i]] = p w^(p-1)
    del t_lat, t_pri, t_perm, w
    glat = lat.clone()
    total = F.swd_segments(glat, pri.clone(), p)
    assert abs(total.item() - total_ref) <= 1e-6 * total_ref, (total.item(), total_ref)
    assert dev_rel(glat, g_ref) <= 1e-6

    # 4. fused entry point vs the stages, and the z gradient vs float64
    tot_f, grad_f = F.swd_sweep(z, prior, proj, p)
    assert abs(tot_f.item() - total.item()) <= 1e-12 * total_ref
    assert dev_rel(grad_f, F.swd_project_back(glat, proj)) <= 1e-6
    g64 = g_ref.double().t() @ proj.double()
    e = dev_rel(grad_f, g64)
    print("full-size swd: sum %.9e vs fp64 %.9e; grad rel err vs fp64 %.2e" % (tot_f.item(), total_ref, e))
    assert e <= 1e-5
    del g64, g_ref

    # 5. idempotence and batch-permutation invariance
    tot0, g0 = F.swd_sweep(z, z.clone(), proj, p)
    assert tot0.item() == 0.0 and not g0.any()
    rp = torch.randperm(n, device=DEV)
    lat_p = F.swd_project(z[rp], proj)
    assert torch.equal(lat_p, lat[:, rp]), "a row's projections depend on its position in the batch"
    tot_p = F.swd_segments(lat_p, pri.clone(), p)                # lat_p now holds d sum / d lat of the permuted batch
    assert abs(tot_p.item() - total.item()) <= 1e-9 * total_ref    # same sorted values, other grouping of the fp32 partials
    # only samples whose key is bit-identical to another sample's may trade ranks (the sort is stable in batch order)
    tied_p, want = tied[:, rp], glat[:, rp]
    frac = tied.float().mean().item()
    print("full-size swd: %.2f %% of the (projection, sample) keys are exact ties" % (100 * frac))
    assert frac < 0.2
    assert torch.equal(lat_p[~tied_p], want[~tied_p]), "an untied sample's d/dlat changed under a batch permutation"
    del lat_p, want, tied, tied_p, glat, lat, pri
    tot_q, grad_p = F.swd_sweep(z[rp], prior, proj, p)
    assert abs(tot_q.item() - tot_f.item()) <= 1e-9 * total_ref
    e = dev_rel(grad_p, grad_f[rp])
    print("full-size swd: gradient under a batch permutation differs by %.2e of its maximum (tied keys in the tails, "
          "where neighbouring prior order statistics are far apart)" % e)
    assert e <= 1e-3


@pytest.mark.parametrize("p", [1.0, 3.0, 2.5])
def test_swd_other_wasserstein_degrees(vk, p):
    """swae.py:178 `w.pow(p)` for degrees other than the shipped p = 2: odd degrees keep the sign of the sorted difference
    (the 'distance' can be negative, as in the reference), p = 1 has the constant gradient 1, and a fractional degree
    turns negative differences into NaN exactly as torch.pow does."""
    F, _ = vk
    gen = torch.Generator().manual_seed(31 + int(10 * p))
    n, D, S = 777, 24, 13
    z = 1.5 * torch.randn(n, D, generator=gen) + 0.1
    pz = torch.randn(n, D, generator=gen)
    th = O.random_projections(D, S)
    zc = leaf(z)
    ref = O.swd(zc.double(), pz.double(), th.double(), p, 0.01)
    zg = leaf(z.to(DEV))
    s = F.swd(zg, pz.to(DEV), th.to(DEV), p, 0.01)
    if p == 2.5:
        assert torch.isnan(ref) and torch.isnan(s)
        return
    ref.backward()
    s.backward()
    assert_close(s, ref, TOL, "swd p=%g" % p)
    clear = rank_clear_rows(z, th)
    assert clear.float().mean() > 0.9
    assert_close(zg.grad.cpu()[clear], zc.grad[clear], TOL, "d swd / dz p=%g" % p)
