"""Two-GPU NCCL parity of vaekit.dist: the row-sharded MMD and the projection-sharded SWD of the
GLOBAL batch must equal the single-GPU kernels run on the gathered batch with the same shared prior, and both must
match the CPU oracle.  Needs >= 2 visible GPUs (skipped otherwise); run with `gpurun --gpus 2`."""
import os
import socket
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "pytorch-vae_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

pytestmark = pytest.mark.gpu

N_LOCAL, D, S = 384, 48, 10
SEED = 4242


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _local_z(rank, dev):
    g = torch.Generator(device="cpu").manual_seed(900 + rank)
    return (torch.randn(N_LOCAL, D, generator=g) * 1.2 + 0.25).to(dev)


def _worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import vae_oracle as O
        from vaekit import dist as vd
        res = {}
        for kt in ("rbf", "imq"):
            torch.manual_seed(SEED)                                   # rank 0's generator state is the shared one
            z = _local_z(rank, dev).requires_grad_(True)
            mmd, terms = vd.sharded_mmd(z, 3.0, kt, 2.0)
            mmd.backward()
            res[kt] = (mmd.detach().cpu(), terms.cpu(), z.grad.cpu())
        # Beta-TC latent terms, row-sharded
        gtc = torch.Generator(device="cpu").manual_seed(700 + rank)
        mu = torch.randn(N_LOCAL, D, generator=gtc).to(dev).requires_grad_(True)
        lv = (0.3 * torch.randn(N_LOCAL, D, generator=gtc)).to(dev).requires_grad_(True)
        eps = torch.randn(N_LOCAL, D, generator=gtc).to(dev)
        ztc = mu + torch.exp(0.5 * lv) * eps
        tc_loss, tc_terms = vd.sharded_betatc_terms(ztc, mu, lv, 0.005, 1.0, 6.0, 0.7)
        (tc_loss * 1.3).backward()
        res["tc"] = (tc_loss.detach().cpu(), tc_terms.cpu(), mu.grad.cpu(), lv.grad.cpu(),
                     mu.detach().cpu(), lv.detach().cpu(), eps.cpu())
        proj = O.random_projections(D, S).to(dev) if rank == 0 else torch.zeros(S, D, device=dev)
        vd.shared_tensor_(proj)
        for key, exchange, waves in (("swd", "all_to_all", None), ("swd_w1", "all_to_all", 1), ("swd_w3", "all_to_all", 3),
                                     ("swd_ag", "all_gather", None), ("swd_peer", "peer", None), ("swd_peer2", "peer", None)):
            torch.manual_seed(SEED)
            z = _local_z(rank, dev).requires_grad_(True)
            swd = vd.sharded_swd(z, proj, 2.0, 7.0, exchange=exchange, waves=waves)
            swd.backward()
            res[key] = (swd.detach().cpu(), z.grad.cpu())
        res["proj"] = proj.cpu()
        torch.save(res, out + ".r%d" % rank)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_sharded_ops_match_single_gpu_and_oracle(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    import vae_oracle as O
    from vaekit import functional as F
    world = 2
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    parts = [torch.load(out + ".r%d" % r) for r in range(world)]
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    z_all = torch.cat([_local_z(r, dev) for r in range(world)])
    N = z_all.shape[0]

    def shared_prior():
        torch.manual_seed(SEED)
        return F.philox_normal(z_all.shape, dev, rng=F.next_philox(dev, z_all.numel()))

    for kt in ("rbf", "imq"):
        prior = shared_prior()
        zs = z_all.clone().requires_grad_(True)
        one, terms = F.mmd(zs, prior, 3.0, kt, 2.0)
        one.backward()
        for r in range(world):
            got, gterms, _ = parts[r][kt]
            assert abs(got.item() - one.item()) <= 1e-6 * abs(one.item()) + 1e-9, (kt, r)
            assert torch.allclose(gterms, terms.cpu(), rtol=1e-6, atol=0)
        g = torch.cat([parts[r][kt][2] for r in range(world)])
        assert ((g - zs.grad.cpu()).abs().max() / zs.grad.abs().max().cpu()).item() < 1e-6, kt
        zo = z_all.cpu().double().requires_grad_(True)
        ref = O.mmd(zo, prior.cpu().double(), 3.0, kt, 2.0)
        ref.backward()
        ref_terms = O.mmd_terms(z_all.cpu().double(), prior.cpu().double(), kt, 2.0)        # (pp, zz, pz)
        for a, b in zip(terms.tolist(), [t.item() for t in ref_terms]):
            assert abs(a - b) <= 1e-5 * abs(b), (kt, a, b)
        scale = 3.0 * max(abs(t.item()) for t in ref_terms)       # the statistic is a difference of these sums
        assert abs(one.item() - ref.item()) <= 1e-5 * scale, kt
        assert ((g.double() - zo.grad).abs().max() / zo.grad.abs().max()).item() < 1e-5, kt

    # Beta-TC: sharded == single-GPU kernels on the gathered batch (gradients flow into mu / log_var through z too)
    mu = torch.cat([parts[r]["tc"][4] for r in range(world)]).to(dev).requires_grad_(True)
    lv = torch.cat([parts[r]["tc"][5] for r in range(world)]).to(dev).requires_grad_(True)
    eps = torch.cat([parts[r]["tc"][6] for r in range(world)]).to(dev)
    one, terms = F.betatc_terms(mu + torch.exp(0.5 * lv) * eps, mu, lv, 0.005, 1.0, 6.0, 0.7)
    (one * 1.3).backward()
    for r in range(world):
        assert abs(parts[r]["tc"][0].item() - one.item()) <= 2e-6 * abs(one.item()), r
        assert torch.allclose(parts[r]["tc"][1], terms.cpu(), rtol=2e-6, atol=1e-7)
    for k, ref in ((2, mu.grad), (3, lv.grad)):
        g = torch.cat([parts[r]["tc"][k] for r in range(world)])
        assert ((g - ref.cpu()).abs().max() / ref.abs().max().cpu()).item() < 2e-6, k
    mo, lo = mu.detach().cpu().double().requires_grad_(True), lv.detach().cpu().double().requires_grad_(True)
    zo = mo + torch.exp(0.5 * lo) * eps.cpu().double()
    ref = O.betatc_loss(torch.zeros(N, 1, dtype=torch.float64), torch.zeros(N, 1, dtype=torch.float64), mo, lo, zo, 0.005,
                        alpha=1.0, beta=6.0, gamma=1.0, anneal_rate=0.7)
    assert abs(one.item() - ref["loss"].item()) <= 1e-5 * abs(ref["loss"].item())

    prior = shared_prior()
    proj = parts[0]["proj"].to(dev)
    assert torch.equal(parts[0]["proj"], parts[1]["proj"])
    zs = z_all.clone().requires_grad_(True)
    one = F.swd(zs, prior, proj, 2.0, 7.0)
    one.backward()
    for key in ("swd", "swd_w1", "swd_w3", "swd_ag", "swd_peer", "swd_peer2"):     # peer twice: both buffer parities                               # all-to-all of projected values / all-gather of z
        for r in range(world):
            assert abs(parts[r][key][0].item() - one.item()) <= 1e-6 * abs(one.item()), (key, r)
        g = torch.cat([parts[r][key][1] for r in range(world)])
        assert ((g - zs.grad.cpu()).abs().max() / zs.grad.abs().max().cpu()).item() < 1e-5, key
    zo = z_all.cpu().requires_grad_(True)
    ref = O.swd(zo, prior.cpu(), proj.cpu(), 2.0, 7.0)
    ref.backward()
    assert abs(one.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert N == world * N_LOCAL
