"""Multi-GPU decompositions of the two global-batch statistics (one process per GPU, torch.distributed).

The reference never calls a collective itself: under Lightning DDP (run.py:86) every rank evaluates MMD / SWD on
its *local* batch.  The sharded modes below are new capability: the statistic of the GLOBAL batch,
with the quadratic / sort work split across ranks.

* MMD (wae_mmd.py:193-204): all-gather z -> every rank holds [N, D]; rank r owns a 128-aligned row block and sweeps
  its rows against all N columns (vk_mmd_fwd row range).  prior_z is regenerated locally from the shared Philox
  (seed, offset) -- no traffic.  Each rank has complete rows, so the gradient of its own z rows is local: no
  reduce-scatter in backward.  One all-reduce of three float64 sums.
* SWD (swae.py:151-179): every rank projects its own rows (z and its slice of the shared prior stream) on all S
  projections; ONE all-to-all of the projected values gives rank r all N values of its S / world projections, which it
  sorts; d/dlat goes back through a second all-to-all and the gradient GEMM runs on the rank's own rows; one
  all-reduce of the float64 loss sum.  (The older all-gather-z / reduce-scatter-gradient form is kept as
  exchange="all_gather".)

* BetaTC (betatc_vae.py:165-205): all-gather z, mu, log_var; rank r sweeps its anchor rows against all columns; the four
  batch means are all-reduced, log_q_z and the per-dimension LSE rows all-gathered for the backward's column sweep; the
  gradient of the rank's own rows is complete locally.

The batch-local ops (K1-K3, K7) need no collective at all: shard the batch, which is what DDP already does.
The plumbing helpers are backend-agnostic (NCCL on GPUs; gloo in the CPU tests).
"""
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

ROW_ALIGN = 128      # row-block granularity of the Gram kernel (vk_mmd_fwd)


# ---------------------------------------------------------------------------
# partition plans (pure host logic)
# ---------------------------------------------------------------------------
def plan_rows(n_local: int, world: int, rank: int, align: int = ROW_ALIGN) -> Tuple[int, int, int]:
    """(N, row_begin, row_end) of the rows rank `rank` owns when every rank contributes `n_local` rows.
    The Gram kernel needs 128-aligned row blocks, so n_local must be a multiple of `align` unless world == 1."""
    if world < 1 or not (0 <= rank < world) or n_local < 1:
        raise ValueError("bad partition request: n_local=%d world=%d rank=%d" % (n_local, world, rank))
    if world > 1 and n_local % align != 0:
        raise ValueError("row-sharded MMD needs a per-rank batch that is a multiple of %d (got %d)" % (align, n_local))
    return n_local * world, rank * n_local, (rank + 1) * n_local


def plan_projections(num_projections: int, world: int, rank: int) -> Tuple[int, int]:
    """[s0, s1) of the projections rank `rank` owns: contiguous, sizes differ by at most one, empty slices allowed."""
    if world < 1 or not (0 <= rank < world) or num_projections < 0:
        raise ValueError("bad partition request: S=%d world=%d rank=%d" % (num_projections, world, rank))
    base, extra = divmod(num_projections, world)
    s0 = rank * base + min(rank, extra)
    return s0, s0 + base + (1 if rank < extra else 0)


# ---------------------------------------------------------------------------
# collectives (thin wrappers; identity when no process group is active)
# ---------------------------------------------------------------------------
def _active(group) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def world_rank(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def gather_rows(x_local: torch.Tensor, group=None) -> torch.Tensor:
    """[n_local, D] on every rank -> [world * n_local, D] on every rank (rank-major)."""
    if not _active(group):
        return x_local.contiguous()
    world = dist.get_world_size(group)
    out = torch.empty((world * x_local.shape[0],) + tuple(x_local.shape[1:]), dtype=x_local.dtype, device=x_local.device)
    dist.all_gather_into_tensor(out, x_local.contiguous(), group=group)
    return out


def gather_rows_async(x_local: torch.Tensor, group=None):
    """gather_rows started asynchronously: (output, work handle or None).  `handle.wait()` orders the CURRENT stream after
    the collective; kernels enqueued before the wait run under it."""
    if not _active(group):
        return x_local.contiguous(), None
    world = dist.get_world_size(group)
    out = torch.empty((world * x_local.shape[0],) + tuple(x_local.shape[1:]), dtype=x_local.dtype, device=x_local.device)
    work = dist.all_gather_into_tensor(out, x_local.contiguous(), group=group, async_op=True)
    return out, work


def reduce_sum_(t: torch.Tensor, group=None) -> torch.Tensor:
    if _active(group):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def scatter_rows_sum(x_all: torch.Tensor, group=None) -> torch.Tensor:
    """Sum of every rank's [world * n_local, D] tensor, each rank keeping its own row block."""
    if not _active(group):
        return x_all
    world = dist.get_world_size(group)
    out = torch.empty((x_all.shape[0] // world,) + tuple(x_all.shape[1:]), dtype=x_all.dtype, device=x_all.device)
    dist.reduce_scatter_tensor(out, x_all.contiguous(), op=dist.ReduceOp.SUM, group=group)
    return out


def shared_rng(device, numel: int, group=None) -> Tuple[int, int]:
    """Philox (seed, offset) for `numel` normals, identical on every rank (rank 0's generator state wins)."""
    from . import functional as F
    seed, offset = F.next_philox(device, numel)
    if _active(group):
        t = torch.tensor([seed - (1 << 63), offset], dtype=torch.int64, device=device)     # seed as a signed payload
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        seed, offset = int(t[0].item()) + (1 << 63), int(t[1].item())
    return seed, offset


def shared_rng_key(device, numel: int, group=None) -> torch.Tensor:
    """The same reservation as shared_rng, as an int64 CUDA tensor [seed, offset] (bit patterns) that is identical on
    every rank after an asynchronous broadcast of rank 0's values: no host synchronisation in the step
    (vk_philox_normal_dev reads the key on the device)."""
    from . import functional as F
    seed, offset = F.next_philox(device, numel)
    signed = seed - (1 << 64) if seed >= (1 << 63) else seed
    # pinned staging + non-blocking copy: a pageable host->device copy would synchronise the stream every step
    host = torch.tensor([signed, offset], dtype=torch.int64).pin_memory() if device.type == "cuda" else \
        torch.tensor([signed, offset], dtype=torch.int64)
    key = host.to(device, non_blocking=True)
    if _active(group):
        dist.broadcast(key, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return key


def shared_projections(model, latent_dim: int, num_projections: int, device, group=None) -> torch.Tensor:
    """The reference's CPU-generated projection directions (swae.py:129-148 via model.get_random_projections), moved to
    `device` through pinned memory without a host synchronisation and made identical on every rank (rank 0's draw)."""
    proj = model.get_random_projections(latent_dim, num_samples=num_projections)
    if torch.device(device).type == "cuda":
        proj = proj.contiguous().pin_memory().to(device, non_blocking=True)
    return shared_tensor_(proj, group)


def shared_tensor_(t: torch.Tensor, group=None) -> torch.Tensor:
    if _active(group):
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return t


# ---------------------------------------------------------------------------
# sharded ops (CUDA)
# ---------------------------------------------------------------------------
class _ShardedMmd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z_local, kernel_type, z_var, reg_weight, passes, group):
        from . import functional as F
        world, rank = world_rank(group)
        N, r0, r1 = plan_rows(z_local.shape[0], world, rank)
        # the gather is started first and the prior (a pure function of the shared key, no traffic) is drawn under it
        z_all, pending = gather_rows_async(z_local, group)
        prior_all = F.philox_normal_dev((N, z_local.shape[1]), shared_rng_key(z_local.device, N * z_local.shape[1], group))
        if pending is not None:
            pending.wait()
        sums, grad = F.mmd_sweep(z_all, prior_all, kernel_type, z_var, r0, r1, passes)
        reduce_sum_(sums, group)
        out = F.mmd_combine(sums, N, kernel_type, reg_weight)
        ctx.save_for_backward(grad)
        ctx.factor = 4.0 * reg_weight * (1.0 / (float(N) * float(N)) if kernel_type == 0 else 1.0)
        ctx.set_materialize_grads(False)
        terms = out[1:].detach()
        ctx.mark_non_differentiable(terms)
        return out[0], terms

    @staticmethod
    def backward(ctx, g, _g_terms):
        from . import functional as F
        if g is None:
            return (None,) * 6
        return F.scale_by_scalar(ctx.saved_tensors[0], g.contiguous(), ctx.factor), None, None, None, None, None


def sharded_mmd(z_local, reg_weight, kernel_type, z_var, passes=0, group=None):
    """MMD of the GLOBAL batch (all ranks' z rows) against a shared prior sample; returns (mmd, terms[3]) identical
    on every rank; backward gives d mmd / d z_local."""
    from . import functional as F
    if kernel_type not in F.KERNEL_TYPES:
        raise ValueError('Undefined kernel type.')
    return _ShardedMmd.apply(z_local, F.KERNEL_TYPES[kernel_type], float(z_var), float(reg_weight), int(passes), group)


class _ShardedBetaTc(torch.autograd.Function):
    """betatc_vae.py:165-205 (latent terms) on the GLOBAL batch: all-gather z, mu, log_var; rank r sweeps its anchor rows
    against all B columns (vk_tc_fwd_rows); the four batch means are all-reduced; log_q_z / the per-dimension LSE rows are
    all-gathered because the column sweep of the backward (g_mu, g_lv of the rank's own rows as PARTNERS) sums over every
    anchor.  Every rank ends with the complete gradient of its own rows: no reduce-scatter."""

    @staticmethod
    def forward(ctx, z_local, mu_local, lv_local, M_N, alpha, beta, gamma_anneal, group):
        from . import functional as F
        world, rank = world_rank(group)
        n = z_local.shape[0]
        r0, r1 = rank * n, (rank + 1) * n
        z_all, mu_all, lv_all = (gather_rows(t, group) for t in (z_local, mu_local, lv_local))
        B = z_all.shape[0]
        out, lqz, lse = F.betatc_rows_fwd(z_all, mu_all, lv_all, r0, r1, M_N, alpha, beta, gamma_anneal)
        reduce_sum_(out, group)
        lqz_all = gather_rows(lqz.view(2, B)[:, r0:r1].t().contiguous(), group).t().contiguous().view(2 * B)
        lse_all = gather_rows(lse[r0:r1], group)
        ctx.save_for_backward(z_all, mu_all, lv_all, lqz_all, lse_all)
        ctx.cfg = (r0, r1, M_N, alpha, beta, gamma_anneal)
        ctx.set_materialize_grads(False)
        terms = out[1:].detach()
        ctx.mark_non_differentiable(terms)
        return out[0], terms

    @staticmethod
    def backward(ctx, g, _g_terms):
        from . import functional as F
        if g is None:
            return (None,) * 8
        z_all, mu_all, lv_all, lqz_all, lse_all = ctx.saved_tensors
        r0, r1, M_N, alpha, beta, gamma_anneal = ctx.cfg
        g_z, g_mu, g_lv = F.betatc_rows_bwd(z_all, mu_all, lv_all, lqz_all, lse_all, r0, r1, M_N, alpha, beta, gamma_anneal, g)
        return g_z, g_mu, g_lv, None, None, None, None, None


def sharded_betatc_terms(z_local, mu_local, log_var_local, M_N, alpha, beta, gamma_anneal, group=None):
    """(alpha*mi + beta*tc + gamma_anneal*kld, terms[3] = {kld, tc, mi}) of the GLOBAL batch, identical on every rank;
    backward gives the gradient w.r.t. the rank's own rows.  M_N refers to the global batch (dataset_size = B_global / M_N,
    betatc_vae.py:176).  Every rank must contribute the same number of rows; latent dim <= 128."""
    return _ShardedBetaTc.apply(z_local, mu_local, log_var_local, float(M_N), float(alpha), float(beta),
                                float(gamma_anneal), group)


class _ShardedSwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z_local, proj, p, reg_weight, group):
        from . import functional as F
        world, rank = world_rank(group)
        z_all = gather_rows(z_local, group)
        N, S = z_all.shape[0], proj.shape[0]
        prior_all = F.philox_normal_dev(z_all.shape, shared_rng_key(z_local.device, z_all.numel(), group))
        s0, s1 = plan_projections(S, world, rank)
        if s1 > s0:
            total, grad_all = F.swd_sweep(z_all, prior_all, proj[s0:s1], p, want_grad=True)
        else:
            total = torch.zeros(1, dtype=torch.float64, device=z_all.device)
            grad_all = torch.zeros_like(z_all)
        reduce_sum_(total, group)
        grad_local = scatter_rows_sum(grad_all, group)
        ctx.save_for_backward(grad_local)
        count = float(N) * float(S)
        ctx.factor = reg_weight / count
        ctx.set_materialize_grads(False)
        return F.mean_scale(total, count, reg_weight)

    @staticmethod
    def backward(ctx, g):
        from . import functional as F
        if g is None:
            return (None,) * 5
        return F.scale_by_scalar(ctx.saved_tensors[0], g.contiguous(), ctx.factor), None, None, None, None


# ---- exchange of projected values (all-to-all): every stage of the SWD shrinks with the world size ----------------
def exchange_segments(local: torch.Tensor, group=None) -> torch.Tensor:
    """[S, C, n] on every rank (all S projections of the rank's own n rows, C channels) -> [S_r, C, world * n]: all rows
    of the projections [s0, s1) this rank owns (plan_projections), rows in rank-major (= global) order."""
    if not _active(group):
        return local.contiguous()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    S, C, n = local.shape
    splits = [b - a for a, b in (plan_projections(S, world, r) for r in range(world))]
    mine = splits[rank]
    recv = torch.empty((world * mine, C, n), dtype=local.dtype, device=local.device)
    dist.all_to_all_single(recv, local.contiguous(), output_split_sizes=[mine] * world, input_split_sizes=splits,
                           group=group)
    return recv.view(world, mine, C, n).permute(1, 2, 0, 3).reshape(mine, C, world * n)


def return_segments(seg: torch.Tensor, num_projections: int, group=None) -> torch.Tensor:
    """Inverse of exchange_segments for one channel: [S_r, world * n] (this rank's projections, all rows) -> [S, n]
    (all projections, this rank's rows)."""
    if not _active(group):
        return seg.contiguous()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    mine, N = seg.shape
    n = N // world
    splits = [b - a for a, b in (plan_projections(num_projections, world, r) for r in range(world))]
    send = seg.view(mine, world, n).permute(1, 0, 2).contiguous().view(world * mine, n)
    recv = torch.empty((num_projections, n), dtype=seg.dtype, device=seg.device)
    dist.all_to_all_single(recv, send, output_split_sizes=splits, input_split_sizes=[mine] * world, group=group)
    return recv


def local_prior_rows(n_local: int, D: int, key: torch.Tensor, rank: int) -> torch.Tensor:
    """Rows [rank * n_local, (rank + 1) * n_local) of the [N, D] prior sample of the shared Philox stream `key`
    (shared_rng_key): noise is a pure function of (seed, offset, element index) with one counter per four elements, so
    a rank's row block is the same stream started rank * n_local * D / 4 counters later -- nothing is communicated."""
    from . import functional as F
    first = rank * n_local * D
    if first % 4 != 0:
        raise ValueError("sharded prior needs n_local * D to be a multiple of 4 (got %d x %d)" % (n_local, D))
    return F.philox_normal_dev((n_local, D), key, quad_shift=first // 4)


def plan_waves(num_projections: int, world: int, waves: int) -> List[List[Tuple[int, int]]]:
    """Pipelining plan of the all-to-all SWD: wave q holds, for every rank r, the q-th piece [a, b) of the projections
    rank r owns (plan_projections split again into `waves` pieces).  Returns waves x world ranges."""
    out = []
    for q in range(waves):
        per_rank = []
        for r in range(world):
            s0, s1 = plan_projections(num_projections, world, r)
            c0, c1 = plan_projections(s1 - s0, waves, q)
            per_rank.append((s0 + c0, s0 + c1))
        out.append(per_rank)
    return out


def _all_to_all(out, inp, out_splits, in_splits, group, async_op):
    if not _active(group):
        out.copy_(inp)
        return None
    return dist.all_to_all_single(out, inp, output_split_sizes=out_splits, input_split_sizes=in_splits, group=group,
                                  async_op=async_op)


class _ShardedSwdA2A(torch.autograd.Function):
    """swae.py:151-179 on the global batch with every stage sharded: rank r projects ITS rows (z and the locally
    regenerated prior rows) on all S projections, an all-to-all hands rank r all N values of its S / world projections,
    it sorts them and forms the loss partial and d/dlat, a second all-to-all returns d/dlat to the row owners, and the
    gradient GEMM runs on the rank's own rows.  No [N, D] tensor is gathered or reduced.  The projections travel in
    `waves` pieces so the exchange of wave q + 1 (and the return of wave q - 1) overlaps the sort of wave q."""

    @staticmethod
    def forward(ctx, z_local, proj, p, reg_weight, group, waves):
        from . import functional as F
        world, rank = world_rank(group)
        n, D = z_local.shape
        S = proj.shape[0]
        N = n * world
        dev = z_local.device
        key = shared_rng_key(dev, N * D, group)
        prior_local = local_prior_rows(n, D, key, rank)
        plan = plan_waves(S, world, waves)
        order = [i for per_rank in plan for a, b in per_rank for i in range(a, b)]      # projections, wave-major
        proj_w = proj[torch.tensor(order, dtype=torch.long, device=dev)] if waves > 1 else proj
        # ---- stage 1: project the rank's rows wave by wave and start every exchange
        both_all = torch.empty((S, 2, n), dtype=torch.float32, device=dev)       # wave-major projection order
        F.swd_project(z_local, proj_w, out=both_all[:, 0, :])
        F.swd_project(prior_local, proj_w, out=both_all[:, 1, :])
        pending, row0 = [], 0
        for per_rank in plan:
            splits = [b - a for a, b in per_rank]
            Sq, mine = sum(splits), splits[rank]
            both = both_all[row0:row0 + Sq]
            recv = torch.empty((world * mine, 2, n), dtype=torch.float32, device=dev)
            work = _all_to_all(recv, both, [mine] * world, splits, group, True)
            pending.append((work, recv, both, splits, mine, row0))
            row0 += Sq
        # ---- stage 2: per wave, sort + loss + d/dlat on the rank's projections, and send d/dlat back
        g_all = torch.empty((S, n), dtype=torch.float32, device=dev)
        total = torch.zeros(1, dtype=torch.float64, device=dev)
        returning = []
        for work, recv, both, splits, mine, r0 in pending:
            if work is not None:
                work.wait()
            Sq = sum(splits)
            if mine > 0:
                v = recv.view(world, mine, 2, n)
                lat = v[:, :, 0, :].permute(1, 0, 2).reshape(mine, N)              # rows in rank-major = global order
                pri = v[:, :, 1, :].permute(1, 0, 2).reshape(mine, N)
                total += F.swd_segments(lat, pri, p, want_grad=True)                # lat <- d sum / d lat, sample order
                send = lat.view(mine, world, n).permute(1, 0, 2).contiguous().view(world * mine, n)
            else:
                send = torch.empty((0, n), dtype=torch.float32, device=dev)
            back = _all_to_all(g_all[r0:r0 + Sq], send, splits, [mine] * world, group, True)
            returning.append((back, send))
        reduce_sum_(total, group)
        for back, _ in returning:
            if back is not None:
                back.wait()
        grad_local = F.swd_project_back(g_all, proj_w)
        ctx.save_for_backward(grad_local)
        count = float(N) * float(S)
        ctx.factor = reg_weight / count
        ctx.set_materialize_grads(False)
        return F.mean_scale(total, count, reg_weight)

    @staticmethod
    def backward(ctx, g):
        from . import functional as F
        if g is None:
            return (None,) * 6
        return F.scale_by_scalar(ctx.saved_tensors[0], g.contiguous(), ctx.factor), None, None, None, None, None


# ---- exchange fused into the GEMMs over peer-mapped memory (NVLink / NVSwitch) ---------------------------------------
class _PeerSegments:
    """Symmetric (peer-mapped) segment buffers of one (group, n_local, S) shape: on every rank a [2 parities][2 channels
    (z, prior)][S_max][N] fp32 block, plus the column tables that tell the GEMMs where each projection lives:
      store[parity][channel][s] = address, in the memory of the rank that OWNS projection s, of that rank's
                                  segment row for s at column rank * n (where this rank's n values go);
      load[parity][s]           = address of the same spot in the z channel (after the sort it holds d/dlat)."""

    def __init__(self, group, device, n, S):
        import torch.distributed._symmetric_memory as symm
        world, rank = world_rank(group)
        self.world, self.rank, self.n, self.S, self.N = world, rank, n, S, n * world
        self.parts = [plan_projections(S, world, r) for r in range(world)]
        self.s_max = max(b - a for a, b in self.parts)
        self.mine = self.parts[rank][1] - self.parts[rank][0]
        shape = (2, 2, self.s_max, self.N)
        self.buf = symm.empty(shape, dtype=torch.float32, device=device)
        self.hdl = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        ptrs = self.hdl.buffer_ptrs
        self.store, self.load = [], []
        for parity in range(2):
            per_chan = []
            for chan in range(2):
                addr = []
                for r, (a, b) in enumerate(self.parts):
                    for j in range(b - a):
                        off = ((parity * 2 + chan) * self.s_max + j) * self.N + rank * n
                        addr.append(ptrs[r] + 4 * off)
                per_chan.append(torch.tensor(addr, dtype=torch.int64, device=device))
            self.store.append(per_chan)
            self.load.append(per_chan[0])
        self.parity = 0

    def segments(self, parity):
        """This rank's (lat, pri) [S_r, N] views of the local block."""
        return self.buf[parity, 0, :self.mine], self.buf[parity, 1, :self.mine]


_peer_cache = {}


def _peer_segments(group, device, n, S) -> "_PeerSegments":
    key = (id(group) if group is not None else 0, device.index, n, S)
    if key not in _peer_cache:
        _peer_cache[key] = _PeerSegments(group, device, n, S)
    return _peer_cache[key]


class _ShardedSwdPeer(torch.autograd.Function):
    """swae.py:151-179 on the global batch with both exchanges FUSED into the tcgen05 GEMMs over peer memory: the
    projection GEMM's epilogue stores each projected value straight into the segment buffer of the rank that sorts that
    projection (vk_gemm_nt_tab column table), and the gradient GEMM's producer warps read d/dlat columns straight from
    their owners.  No NCCL transfer, no staging copy, no permute; two device-side barriers per step, buffers
    double-buffered by step parity so a rank never writes a block a slower peer is still reading."""

    @staticmethod
    def forward(ctx, z_local, proj, p, reg_weight, group):
        from . import functional as F
        world, rank = world_rank(group)
        n, D = z_local.shape
        S = proj.shape[0]
        N = n * world
        ps = _peer_segments(group, z_local.device, n, S)
        parity, ps.parity = ps.parity, ps.parity ^ 1
        key = shared_rng_key(z_local.device, N * D, group)
        prior_local = local_prior_rows(n, D, key, rank)
        F.swd_project_cols(z_local.contiguous(), proj, ps.store[parity][0])
        F.swd_project_cols(prior_local, proj, ps.store[parity][1])
        ps.hdl.barrier(channel=0)                                  # every rank's projections have landed
        lat, pri = ps.segments(parity)
        total = F.swd_segments(lat, pri, p, want_grad=True)        # lat <- d sum / d lat (sample order)
        ps.hdl.barrier(channel=1)                                  # every rank's d/dlat is final
        reduce_sum_(total, group)
        grad_local = F.swd_project_back_cols(ps.load[parity], n, proj)
        ctx.save_for_backward(grad_local)
        count = float(N) * float(S)
        ctx.factor = reg_weight / count
        ctx.set_materialize_grads(False)
        return F.mean_scale(total, count, reg_weight)

    @staticmethod
    def backward(ctx, g):
        from . import functional as F
        if g is None:
            return (None,) * 5
        return F.scale_by_scalar(ctx.saved_tensors[0], g.contiguous(), ctx.factor), None, None, None, None


_peer_ok = {}


def _gemm_fits(J: int, K: int) -> bool:
    """Shape limits of vk_gemm_nt (csrc/gemm3x.cuh g3::fits): tensor-memory columns and the shared-memory B images."""
    Jp, Kp = (J + 15) // 16 * 16, (K + 15) // 16 * 16
    return Jp <= 256 and Kp <= 256 and 3 * (Kp // 2) + Jp <= 512 and 3 * Jp * Kp * 2 + 256 <= 200 * 1024


def _peer_available(group, z_local, S) -> bool:
    """Can this process group map peer memory (torch symmetric memory over NVLink)?  Decided once per group by trying the
    rendezvous; all ranks take the same branch because the rendezvous itself is collective."""
    key = id(group) if group is not None else 0
    if key not in _peer_ok:
        ok = z_local.is_cuda and dist.get_backend(group) == "nccl"
        if ok:
            try:
                _peer_segments(group, z_local.device, z_local.shape[0], S)
            except Exception as e:                                   # no P2P mapping on this box: NCCL all-to-all instead
                import warnings
                warnings.warn("vaekit: peer-mapped memory unavailable (%s); the sharded SWD uses NCCL all-to-all" % (e,))
                ok = False
        _peer_ok[key] = ok
    return _peer_ok[key]


def sharded_swd(z_local, proj, p, reg_weight, group=None, exchange="auto", waves=None):
    """Sliced-Wasserstein distance of the GLOBAL batch; `proj` [S, D] must be identical on every rank
    (use shared_projections / shared_tensor_ on the reference's CPU-generated directions).  exchange = "auto" (default:
    "peer" where the group can map peer memory, else "all_to_all"), "all_to_all" (projected values travel through NCCL;
    every stage shrinks with the world size; `waves` pieces pipeline the exchange under the sorts), "peer" (the same
    decomposition with both exchanges fused into the GEMMs over peer-mapped memory, CUDA + NVLink only) or "all_gather" (z travels; each rank projects all rows on its slice of
    the projections and the [N, D] gradient is reduce-scattered -- kept for shapes the tcgen05 GEMM does not take)."""
    if exchange == "auto":
        S, D = proj.shape
        if not (_gemm_fits(S, D) and _gemm_fits(D, S)):
            exchange = "all_gather"             # shapes the tcgen05 GEMM does not take: vk_swd_fwd has an FFMA GEMM
        else:
            exchange = "peer" if world_rank(group)[0] > 1 and _peer_available(group, z_local, S) else "all_to_all"
    if exchange == "peer" and world_rank(group)[0] > 1:
        return _ShardedSwdPeer.apply(z_local, proj.contiguous(), float(p), float(reg_weight), group)
    if exchange in ("all_to_all", "peer"):
        if waves is None:                       # pipeline depth: pieces of the rank's projection slice
            world, _ = world_rank(group)
            waves = 1 if world == 1 else max(1, min(2, proj.shape[0] // world))
        return _ShardedSwdA2A.apply(z_local, proj, float(p), float(reg_weight), group, int(waves))
    if exchange == "all_gather":
        return _ShardedSwd.apply(z_local, proj, float(p), float(reg_weight), group)
    raise ValueError("Unknown exchange %r" % (exchange,))
