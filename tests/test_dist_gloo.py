"""World-size-2 gloo tests (CPU) of the multi-GPU host logic in vaekit/dist.py: partition plans, the collectives'
plumbing, and that the row-sharded MMD / projection-sharded SWD decompositions add up to the single-process
oracle.  The per-shard arithmetic here is the CPU oracle (the CUDA sweeps are covered by the -m gpu tests, which
emulate the same shards on one GPU)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "pytorch-vae_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def test_partition_plans():
    from vaekit import dist as vd
    assert vd.plan_rows(256, 4, 2) == (1024, 512, 768)
    assert vd.plan_rows(100, 1, 0) == (100, 0, 100)
    with pytest.raises(ValueError):
        vd.plan_rows(100, 2, 0)                  # not 128-aligned
    with pytest.raises(ValueError):
        vd.plan_rows(128, 2, 2)
    for S, w in ((200, 8), (7, 3), (2, 4), (0, 2)):
        parts = [vd.plan_projections(S, w, r) for r in range(w)]
        assert parts[0][0] == 0 and parts[-1][1] == S
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        sizes = [b - a for a, b in parts]
        assert max(sizes) - min(sizes) <= 1
    # wave plan of the pipelined all-to-all SWD: every projection exactly once, each piece inside its owner's slice
    for S, w, q in ((200, 8, 4), (7, 3, 2), (5, 4, 3), (10, 2, 1)):
        plan = vd.plan_waves(S, w, q)
        assert len(plan) == q and all(len(per_rank) == w for per_rank in plan)
        seen = sorted(i for per_rank in plan for a, b in per_rank for i in range(a, b))
        assert seen == list(range(S))
        for r in range(w):
            s0, s1 = vd.plan_projections(S, w, r)
            pieces = [plan[k][r] for k in range(q)]
            assert pieces[0][0] == s0 and pieces[-1][1] == s1 and all(a[1] == b[0] for a, b in zip(pieces, pieces[1:]))
    assert vd._gemm_fits(200, 128) and vd._gemm_fits(128, 200) and not vd._gemm_fits(300, 128) and not vd._gemm_fits(256, 256)
    # without a process group the collectives are identities
    x = torch.arange(6.0).view(3, 2)
    assert torch.equal(vd.gather_rows(x), x) and torch.equal(vd.scatter_rows_sum(x), x)
    assert vd.world_rank() == (1, 0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import vae_oracle as O
        from vaekit import dist as vd
        torch.set_num_threads(1)
        n_local, D, S = 128, 16, 6
        g = torch.Generator().manual_seed(100 + rank)
        z_local = torch.randn(n_local, D, generator=g) + 0.25
        gp = torch.Generator().manual_seed(7)                       # shared "Philox" stand-in: same on every rank
        prior_all = torch.randn(n_local * world, D, generator=gp)
        proj = O.random_projections(D, S) if rank == 0 else torch.zeros(S, D)
        vd.shared_tensor_(proj)

        # ---- plumbing
        z_all = vd.gather_rows(z_local)
        assert z_all.shape == (n_local * world, D) and torch.equal(z_all[rank * n_local:(rank + 1) * n_local], z_local)
        N, r0, r1 = vd.plan_rows(n_local, world, rank)

        # ---- row-sharded MMD: partial kernel sums over this rank's rows, all-reduced
        res = {}
        for kt in ("rbf", "imq"):
            zl = z_all[r0:r1].clone().requires_grad_(True)
            if kt == "rbf":
                k = lambda a, b: torch.exp(-((a[:, None, :] - b[None, :, :]).pow(2).mean(-1) / (2.0 * D * 2.0)))
                sums = torch.stack([k(zl, z_all).sum() * 2, k(zl, prior_all).sum(),
                                    k(prior_all[r0:r1], prior_all).sum()])
            else:
                C = 2 * D * 2.0
                def k(a, b, off):
                    m = C / (1e-7 + C + (a[:, None, :] - b[None, :, :]).pow(2).sum(-1))
                    idx = torch.arange(a.shape[0])
                    m = m.clone()
                    m[idx, idx + off] = 0
                    return m
                sums = torch.stack([k(zl, z_all, r0).sum() * 2, k(zl, prior_all, r0).sum(),
                                    k(prior_all[r0:r1], prior_all, r0).sum()])
            # (the doubling: z_i enters k(z,z) as a row and as a column; z_all is a constant copy here)
            # value: rows-of-z sums counted once
            with torch.no_grad():
                if kt == "rbf":
                    val = torch.stack([k(zl, z_all).sum(), k(zl, prior_all).sum(), k(prior_all[r0:r1], prior_all).sum()]).double()
                else:
                    val = torch.stack([k(zl, z_all, r0).sum(), k(zl, prior_all, r0).sum(),
                                       k(prior_all[r0:r1], prior_all, r0).sum()]).double()
            vd.reduce_sum_(val)
            m = 1.0 / (N * N) if kt == "rbf" else 1.0
            reg = 3.0
            # local gradient of the global mmd w.r.t. this rank's rows: zz term (both roles of z_i) minus 2 * pz term
            (reg * m * (sums[0] - 2 * sums[1])).backward()
            res[kt] = (val * m, zl.grad.clone())

        # ---- projection-sharded SWD: partial sum + partial [N, D] gradient, reduce-scatter
        s0, s1 = vd.plan_projections(S, world, rank)
        za = z_all.clone().requires_grad_(True)
        pm = proj[s0:s1].t()
        w = torch.sort((za @ pm).t(), dim=1)[0] - torch.sort((prior_all @ pm).t(), dim=1)[0]
        tot = w.pow(2).sum()
        tot.backward()
        total = tot.detach().double().view(1)
        vd.reduce_sum_(total)
        grad_local = vd.scatter_rows_sum(za.grad)

        # ---- all-to-all exchange of projected values: [S, C, n] per rank -> [S_r, C, N], and back
        Sx = 7                                                       # uneven split: 4 + 3
        full = torch.arange(Sx * 2 * N, dtype=torch.float32).view(Sx, 2, N)        # the global [S, C, N] table
        mine = full[:, :, rank * n_local:(rank + 1) * n_local].contiguous()
        seg = vd.exchange_segments(mine)
        a0, a1 = vd.plan_projections(Sx, world, rank)
        assert torch.equal(seg, full[a0:a1])
        back = vd.return_segments(seg[:, 1, :].contiguous(), Sx)
        assert torch.equal(back, mine[:, 1, :])
        # the SWD through the exchange: local projections -> segments -> sorted loss / d/dlat -> back -> local gradient
        zl = z_local.clone().requires_grad_(True)
        pl = prior_all[rank * n_local:(rank + 1) * n_local]
        both = torch.stack([zl @ proj.t(), pl @ proj.t()], dim=0).permute(2, 0, 1).contiguous()   # [S, 2, n]
        segs = vd.exchange_segments(both.detach())
        lat = segs[:, 0, :].clone().requires_grad_(True)
        w2 = torch.sort(lat, dim=1)[0] - torch.sort(segs[:, 1, :], dim=1)[0]
        t2 = w2.pow(2).sum()
        t2.backward()
        tot2 = t2.detach().double().view(1)
        vd.reduce_sum_(tot2)
        g_back = vd.return_segments(lat.grad, S)                     # [S, n]
        grad_a2a = g_back.t() @ proj
        assert abs(tot2.item() - total.item()) <= 1e-5 * abs(total.item())
        assert ((grad_a2a - grad_local).abs().max() / grad_local.abs().max()).item() < 1e-5

        if rank == 0:
            torch.save({"z_all": z_all, "prior": prior_all, "proj": proj, "res": res, "swd_total": total,
                        "world": world, "n_local": n_local}, out)
        torch.save({"swd_grad_local": grad_local, "mmd": {k: v[1] for k, v in res.items()}}, out + ".r%d" % rank)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_decompositions_add_up(tmp_path):
    import vae_oracle as O
    world = 2
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    r = torch.load(out)
    z, prior, proj, n_local = r["z_all"], r["prior"], r["proj"], r["n_local"]
    N = z.shape[0]
    parts = [torch.load(out + ".r%d" % k) for k in range(world)]
    for kt in ("rbf", "imq"):
        pp, zz, pz = O.mmd_terms(z, prior, kt, 2.0)
        got = r["res"][kt][0]                                  # {zz, pz, pp}
        for a, b in zip(got.tolist(), (zz.item(), pz.item(), pp.item())):
            assert abs(a - b) <= 1e-5 * abs(b)
        zc = z.clone().requires_grad_(True)
        O.mmd(zc, prior, 3.0, kt, 2.0).backward()
        g = torch.cat([p["mmd"][kt] for p in parts])
        assert ((g - zc.grad).abs().max() / zc.grad.abs().max()).item() < 2e-5
    zc = z.clone().requires_grad_(True)
    ref = O.swd(zc, prior, proj, 2.0, 1.0) * (proj.shape[0] * N)    # un-normalised sum
    ref.backward()
    assert abs(r["swd_total"].item() - ref.item()) <= 1e-5 * abs(ref.item())
    g = torch.cat([p["swd_grad_local"] for p in parts])
    assert ((g - zc.grad).abs().max() / zc.grad.abs().max()).item() < 1e-5
