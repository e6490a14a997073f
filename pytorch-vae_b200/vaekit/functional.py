"""torch.autograd glue over the C-ABI kernels.  Each Function enqueues exactly one
kernel forward and one backward; nothing here computes on the CPU."""
import torch

from . import _lib as L

RECON_MSE_MEAN, RECON_MSE_SUM, RECON_LOGCOSH, RECON_MSE_L1_MEAN, RECON_BCE_MEAN, RECON_BCE_SUM = 0, 1, 2, 3, 4, 5
COMBINE_NONE, COMBINE_LINEAR, COMBINE_ABS, COMBINE_SCALED = 0, 1, 2, 3


# ---------------------------------------------------------------------------
# RNG: (seed, offset) drawn from torch's CUDA generator so torch.manual_seed
# governs the noise; the stream itself is vaekit's own (include/vaekit.h).
# ---------------------------------------------------------------------------
def next_philox(device, numel):
    """Reserve ceil(numel/4) Philox quads on `device`'s default generator."""
    gen = torch.cuda.default_generators[device.index if device.index is not None else torch.cuda.current_device()]
    seed, offset = gen.initial_seed(), gen.get_offset()
    quads = (int(numel) + 3) // 4
    gen.set_offset(offset + ((quads + 3) // 4) * 4)     # torch requires multiples of 4
    return seed & 0xFFFFFFFFFFFFFFFF, offset


# (seed, offset) held ON THE DEVICE: the pair is no launch argument, so a CUDA graph that captured the noise kernels draws
# fresh noise on every replay (vk_reparam_kl_fwd_dev / vk_philox_normal_dev).
_DEVICE_KEYS = {}


def make_rng_key(device):
    """int64 CUDA tensor [2] = (seed, offset) taken from torch's CUDA generator for `device` (torch.manual_seed governs
    it); create it OUTSIDE a graph capture."""
    device = torch.device(device)
    seed, offset = next_philox(device, 0)
    if seed >= 1 << 63:
        seed -= 1 << 64
    return torch.tensor([seed, offset], dtype=torch.int64, device=device)


class device_rng:
    """`with device_rng(key):` -- every noise draw of reparam_kl / philox_normal inside the block reads (seed, offset)
    from `key` on the device and advances key[1] there (two tiny captured torch ops per draw), instead of passing the
    pair by value.  Capture a whole training step in a CUDA graph inside this block and every replay gets new noise."""

    def __init__(self, key):
        if key.dtype != torch.int64 or not key.is_cuda or key.numel() != 2:
            raise RuntimeError("vaekit: rng key must be an int64 CUDA tensor with two elements (make_rng_key)")
        self.key, self.index = key, key.device.index

    def __enter__(self):
        self.prev = _DEVICE_KEYS.get(self.index)
        _DEVICE_KEYS[self.index] = self.key
        return self.key

    def __exit__(self, *exc):
        if self.prev is None:
            _DEVICE_KEYS.pop(self.index, None)
        else:
            _DEVICE_KEYS[self.index] = self.prev
        return False


def _draw_device_key(device, numel):
    """Snapshot of the active device key for one draw of `numel` elements (None outside device_rng); the key advances by
    the same multiple-of-4 quad count next_philox reserves on the host generator."""
    key = _DEVICE_KEYS.get(device.index if device.index is not None else torch.cuda.current_device())
    if key is None:
        return None
    snap = key.clone()
    quads = (int(numel) + 3) // 4
    key[1] += ((quads + 3) // 4) * 4
    return snap


def _row_layout(t):
    """(outer[3], strides[3], D, tensor) for a tensor addressed as rows of D contiguous floats."""
    if t.dim() == 0:
        t = t.reshape(1, 1)
    if t.dim() == 1:
        t = t.unsqueeze(0)
    if t.shape[-1] > 1 and t.stride(-1) != 1:
        t = t.contiguous()
    if t.dim() > 4:                       # more than three outer dims: a dense copy addressed as [rows, D]
        t = t.contiguous().view(-1, t.shape[-1])
    outer = list(t.shape[:-1])
    strides = list(t.stride()[:-1])
    pad = 3 - len(outer)
    return [1] * pad + outer, [0] * pad + strides, t.shape[-1], t


class _ReparamKL(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mu, log_var, seed, offset, want_z, want_kl, want_eps):
        L.require(mu, "mu"); L.require(log_var, "logvar")
        if mu.shape != log_var.shape:
            raise RuntimeError("vaekit: mu %s and logvar %s differ in shape" % (tuple(mu.shape), tuple(log_var.shape)))
        outer, ms, D, mu_v = _row_layout(mu)
        _, ls, _, lv_v = _row_layout(log_var)
        rows = outer[0] * outer[1] * outer[2]
        dev = mu.device
        z = torch.empty(mu.shape, dtype=torch.float32, device=dev) if want_z else None
        kl = torch.empty(mu.shape[:-1], dtype=torch.float32, device=dev) if want_kl else None
        eps = torch.empty(mu.shape, dtype=torch.float32, device=dev) if want_eps else None
        if rows * D > 0:
            with L.on(dev):
                if isinstance(seed, torch.Tensor):          # (seed, offset) on the device
                    L.call("reparam_kl_fwd_dev", L.ptr(mu_v), L.ptr(lv_v), L.i64x3(outer), L.i64x3(ms), L.i64x3(ls), D,
                           L.ptr(seed), L.ptr(z), L.ptr(kl), L.ptr(eps), L.stream_ptr())
                else:
                    L.call("reparam_kl_fwd", L.ptr(mu_v), L.ptr(lv_v), L.i64x3(outer), L.i64x3(ms), L.i64x3(ls), D,
                           seed, offset, L.ptr(z), L.ptr(kl), L.ptr(eps), L.stream_ptr())
        ctx.save_for_backward(mu, log_var)
        ctx.rng = (seed, offset)
        ctx.had_z = want_z
        ctx.set_materialize_grads(False)
        if eps is not None:
            ctx.mark_non_differentiable(eps)
        return z, kl, eps

    @staticmethod
    def backward(ctx, g_z, g_kl, _g_eps):
        mu, log_var = ctx.saved_tensors
        outer, ms, D, mu_v = _row_layout(mu)
        _, ls, _, lv_v = _row_layout(log_var)
        dev = mu.device
        g_mu = torch.empty(mu.shape, dtype=torch.float32, device=dev)
        g_lv = torch.empty(mu.shape, dtype=torch.float32, device=dev)
        if g_z is not None:
            g_z = g_z.contiguous()
        if g_kl is not None:
            g_kl = g_kl.contiguous()
        if g_mu.numel() > 0:
            with L.on(dev):
                if isinstance(ctx.rng[0], torch.Tensor):
                    L.call("reparam_kl_bwd_dev", L.ptr(g_z), L.ptr(g_kl), L.ptr(mu_v), L.ptr(lv_v), L.i64x3(outer),
                           L.i64x3(ms), L.i64x3(ls), D, L.ptr(ctx.rng[0]), L.ptr(g_mu), L.ptr(g_lv), L.stream_ptr())
                else:
                    L.call("reparam_kl_bwd", L.ptr(g_z), L.ptr(g_kl), L.ptr(mu_v), L.ptr(lv_v), L.i64x3(outer),
                           L.i64x3(ms), L.i64x3(ls), D, ctx.rng[0], ctx.rng[1], L.ptr(g_mu), L.ptr(g_lv), L.stream_ptr())
        return g_mu, g_lv, None, None, None, None, None


def reparam_kl(mu, log_var, *, rng=None, want_kl=True, want_eps=False):
    """z = eps*exp(0.5*logvar)+mu with the KL rows as a fused second output.
    Returns (z, kl_rows, eps_or_None); rng = (seed, offset) or None to draw from torch's generator."""
    if rng is None:
        key = _draw_device_key(mu.device, mu.numel()) if mu.is_cuda else None
        rng = (key, 0) if key is not None else next_philox(mu.device, mu.numel())
    elif isinstance(rng, torch.Tensor):
        rng = (rng, 0)
    return _ReparamKL.apply(mu, log_var, rng[0], rng[1], True, want_kl, want_eps)


def kl_rows(mu, log_var):
    """-0.5*sum_d(1+lv-mu^2-exp(lv)) per row (no noise drawn)."""
    return _ReparamKL.apply(mu, log_var, 0, 0, False, True, False)[1]


def philox_normal(shape, device, rng=None):
    """Standard normals from the contract stream (the prior_z draws of WAE / SWAE)."""
    out = torch.empty(shape, dtype=torch.float32, device=device)
    if rng is None:
        key = _draw_device_key(out.device, out.numel()) if out.is_cuda else None
        if key is not None:
            return philox_normal_dev(shape, key)
        rng = next_philox(out.device, out.numel())
    with L.on(out.device):
        L.call("philox_normal", L.ptr(out), out.numel(), rng[0], rng[1], L.stream_ptr())
    return out


def philox_normal_dev(shape, key, quad_shift=0):
    """Standard normals from the contract stream whose (seed, offset) live in the int64 CUDA tensor `key` [2], started
    `quad_shift` counters (4 elements each) later.  No host synchronisation."""
    if key.dtype != torch.int64 or not key.is_cuda or key.numel() != 2:
        raise RuntimeError("vaekit: philox key must be an int64 CUDA tensor with two elements")
    out = torch.empty(shape, dtype=torch.float32, device=key.device)
    with L.on(out.device):
        L.call("philox_normal_dev", L.ptr(out), out.numel(), L.ptr(key), int(quad_shift), L.stream_ptr())
    return out


def philox_raw(n_quads, device, seed, offset):
    out = torch.empty((n_quads, 4), dtype=torch.int32, device=device)
    with L.on(out.device):
        L.call("philox_raw", L.ptr(out), n_quads, seed, offset, L.stream_ptr())
    return out


# ---------------------------------------------------------------------------
# K2
# ---------------------------------------------------------------------------
class _Elbo(torch.autograd.Function):
    @staticmethod
    def forward(ctx, recons, input, kl, recon_mode, alpha, combine_mode, w, C):
        L.require(recons, "recons"); L.require(input, "input")
        if recons.shape != input.shape:
            raise RuntimeError("vaekit: recons %s and input %s differ in shape" % (tuple(recons.shape), tuple(input.shape)))
        if recons.numel() == 0:
            raise RuntimeError("vaekit: empty reconstruction tensor")
        r, x = recons.contiguous(), input.contiguous()
        klc = None
        if kl is not None:
            klc = L.require(kl, "kl_rows").contiguous()
        dev = recons.device
        out = torch.empty(4, dtype=torch.float32, device=dev)
        ws = L.workspace(dev)
        with L.on(dev):
            L.call("elbo_fwd", L.ptr(r), L.ptr(x), r.numel(), L.ptr(klc), 0 if klc is None else klc.numel(),
                                     recon_mode, alpha, combine_mode, w, C, L.ptr(out), L.ptr(ws), ws.numel(),
                                     L.stream_ptr())
        ctx.save_for_backward(r, x, out)
        ctx.cfg = (recon_mode, alpha, combine_mode, w, C, None if klc is None else tuple(kl.shape))
        ctx.set_materialize_grads(False)
        return out[0], out[1], out[2], out[3]

    @staticmethod
    def backward(ctx, g0, g1, g2, g3):
        r, x, out = ctx.saved_tensors
        recon_mode, alpha, combine_mode, w, C, kl_shape = ctx.cfg
        dev = r.device
        need_r, need_x, need_kl = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        if need_x and recon_mode in (RECON_BCE_MEAN, RECON_BCE_SUM):
            raise RuntimeError("vaekit: gradient w.r.t. the target image of the BCE term is not provided")
        g_rec = torch.empty_like(r) if (need_r or need_x) else None
        g_kl = torch.empty(kl_shape, dtype=torch.float32, device=dev) if (need_kl and kl_shape is not None) else None
        gs = [None if g is None else g.contiguous() for g in (g0, g1, g2, g3)]
        rows = 0 if kl_shape is None else max(1, int(torch.Size(kl_shape).numel()))
        with L.on(dev):
            L.call("elbo_bwd", L.ptr(r), L.ptr(x), r.numel(), rows, recon_mode, alpha, combine_mode, w, C,
                                     L.ptr(out), L.ptr(gs[0]), L.ptr(gs[1]), L.ptr(gs[2]), L.ptr(gs[3]),
                                     L.ptr(g_rec), L.ptr(g_kl), L.stream_ptr())
        return (g_rec if need_r else None, (-g_rec) if need_x else None, g_kl, None, None, None, None, None)


def elbo(recons, input, kl=None, *, recon_mode=RECON_MSE_MEAN, alpha=1.0, combine_mode=COMBINE_NONE, w=0.0, C=0.0):
    """Returns (loss, recon_loss, kld, -kld) as 0-dim tensors sharing one 16-byte buffer."""
    return _Elbo.apply(recons, input, kl, int(recon_mode), float(alpha), int(combine_mode), float(w), float(C))


def per_image_mse(recons, input):
    """[B] mean squared error of every image (experiment.py:58-62: mse_loss(reduction='none').mean(dim=[1,2,3])), one
    streaming pass (vk_per_image_mse).  Not differentiable: it is the trainer's bookkeeping value."""
    L.require(recons, "recons"); L.require(input, "input")
    if recons.shape != input.shape or recons.dim() < 2:
        raise RuntimeError("vaekit: recons %s and input %s must be equal-shaped batches" % (tuple(recons.shape), tuple(input.shape)))
    r, x = recons.detach().contiguous(), input.detach().contiguous()
    B = r.shape[0]
    out = torch.empty(B, dtype=torch.float32, device=r.device)
    if r.numel() == 0:
        return out
    with L.on(r.device):
        L.call("per_image_mse", L.ptr(r), L.ptr(x), B, r.numel() // B, L.ptr(out), L.stream_ptr())
    return out


def step_report(loss_dict, per_image=None):
    """Everything the trainer reads back after a step in ONE device-to-host copy: the reference's
    training_step does len(dict) + B separate .item() synchronisations (experiment.py:55, 64-66).  Returns
    ({key: float}, per-image list or None)."""
    keys = list(loss_dict)
    parts = [loss_dict[k].detach().reshape(-1)[:1].float() for k in keys]
    if per_image is not None:
        parts.append(per_image.detach().reshape(-1).float())
    flat = torch.cat(parts).tolist()                         # the step's single synchronisation
    return dict(zip(keys, flat[:len(keys)])), (flat[len(keys):] if per_image is not None else None)


# ---------------------------------------------------------------------------
# K3
# ---------------------------------------------------------------------------
class _IwBound(torch.autograd.Function):
    @staticmethod
    def forward(ctx, recons, input, kl, M_N):
        L.require(recons, "recons"); L.require(input, "input"); L.require(kl, "kl_rows")
        nd = recons.dim() - input.dim()
        if nd not in (1, 2) or recons.shape[0] != input.shape[0] or recons.shape[1 + nd:] != input.shape[1:]:
            raise RuntimeError("vaekit: recons %s is not input %s with 1 or 2 sample dims" %
                               (tuple(recons.shape), tuple(input.shape)))
        if tuple(kl.shape) != tuple(recons.shape[:1 + nd]):
            raise RuntimeError("vaekit: kl rows %s do not match recons sample dims %s" %
                               (tuple(kl.shape), tuple(recons.shape[:1 + nd])))
        B = recons.shape[0]
        G = recons.shape[1] if nd == 2 else 1
        K = recons.shape[nd]
        P = input[0].numel() if B > 0 else 0
        if B * G * K * P == 0:
            raise RuntimeError("vaekit: empty input to the importance-weighted bound")
        r, x, klc = recons.contiguous(), input.contiguous(), kl.contiguous()
        dev = recons.device
        lp = torch.empty(kl.shape, dtype=torch.float32, device=dev)
        out = torch.empty(4, dtype=torch.float32, device=dev)
        ws = L.workspace(dev, 256 + 24 * B)
        with L.on(dev):
            L.call("iw_bound_fwd", L.ptr(r), L.ptr(x), L.ptr(klc), B, G, K, P, M_N, L.ptr(lp), L.ptr(out),
                                         L.ptr(ws), ws.numel(), L.stream_ptr())
        ctx.save_for_backward(r, x, klc, lp)
        ctx.cfg = (B, G, K, P, M_N)
        ctx.set_materialize_grads(False)
        return out[0], out[1], out[2], out[3]

    @staticmethod
    def backward(ctx, g0, g1, g2, g3):
        r, x, klc, lp = ctx.saved_tensors
        B, G, K, P, M_N = ctx.cfg
        if ctx.needs_input_grad[1]:
            raise RuntimeError("vaekit: gradient w.r.t. the input image of the IW bound is not provided")
        g_rec = torch.empty_like(r) if ctx.needs_input_grad[0] else None
        g_kl = torch.empty_like(klc) if ctx.needs_input_grad[2] else None
        gs = [None if g is None else g.contiguous() for g in (g0, g1, g2, g3)]
        with L.on(r.device):
            L.call("iw_bound_bwd", L.ptr(r), L.ptr(x), L.ptr(lp), L.ptr(klc), B, G, K, P, M_N, L.ptr(gs[0]),
                                         L.ptr(gs[1]), L.ptr(gs[2]), L.ptr(gs[3]), L.ptr(g_rec), L.ptr(g_kl),
                                         L.stream_ptr())
        return g_rec, None, g_kl, None


def iw_bound(recons, input, kl, M_N):
    """(loss, mean log_p_x_z, mean kl, -mean kl) of iwae.py:145-160 / miwae.py:148-164."""
    return _IwBound.apply(recons, input, kl, float(M_N))


# ---------------------------------------------------------------------------
# K7
# ---------------------------------------------------------------------------
class _VectorQuantize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, latents, codebook, beta):
        L.require(latents, "latents", dims=(3, 4)); L.require(codebook, "codebook", dims=(2,))
        if latents.shape[1] != codebook.shape[1]:
            raise RuntimeError("vaekit: latents have %d channels but the codebook has embedding_dim %d" %
                               (latents.shape[1], codebook.shape[1]))
        if latents.numel() == 0:
            raise RuntimeError("vaekit: empty latents")
        x, e = latents.contiguous(), codebook.contiguous()
        B, D = x.shape[0], x.shape[1]
        HW = x[0, 0].numel()
        K = e.shape[0]
        dev = x.device
        q = torch.empty_like(x)
        idx = torch.empty(B * HW, dtype=torch.int32, device=dev)
        out = torch.empty(2, dtype=torch.float32, device=dev)
        need = L.ws_bytes("vq", D, K)
        if need < 0:
            raise RuntimeError("vaekit: VectorQuantizer shape D=%d K=%d is not supported" % (D, K))
        ws = L.workspace(dev, need)
        with L.on(dev):
            L.call("vq_fwd", L.ptr(x), L.ptr(e), B, D, HW, K, beta, L.ptr(q), L.ptr(idx), L.ptr(out), L.ptr(ws),
                   ws.numel(), L.stream_ptr())
        ctx.save_for_backward(x, e, idx)
        ctx.cfg = (B, D, HW, K, beta)
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(idx)
        return q, out[0], idx

    @staticmethod
    def backward(ctx, g_q, g_loss, _g_idx):
        x, e, idx = ctx.saved_tensors
        B, D, HW, K, beta = ctx.cfg
        g_x = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        g_e = torch.empty_like(e) if ctx.needs_input_grad[1] else None
        g_q = None if g_q is None else g_q.contiguous()
        g_loss = None if g_loss is None else g_loss.contiguous()
        with L.on(x.device):
            L.call("vq_bwd", L.ptr(g_q), L.ptr(g_loss), L.ptr(x), L.ptr(e), L.ptr(idx), B, D, HW, K, beta,
                   L.ptr(g_x), L.ptr(g_e), L.stream_ptr())
        return g_x, g_e, None


def vector_quantize(latents, codebook, beta):
    """(quantized NCHW with straight-through gradient, vq_loss, indices[int32, B*H*W]) of vq_vae.py:24-55."""
    return _VectorQuantize.apply(latents, codebook, float(beta))


# ---------------------------------------------------------------------------
# K4
# ---------------------------------------------------------------------------
KERNEL_TYPES = {'rbf': 0, 'imq': 1}


MMD_DIAG = None      # set to a list to collect the [4] diagnostics tensor of every vk_mmd_fwd call (tests, bench)


def mmd_sweep(z_all, prior_all, kernel_type, z_var, row_begin, row_end, passes=0):
    """One Gram sweep (vk_mmd_fwd) over rows [row_begin,row_end) of the N x N kernel matrices.
    passes: 0 auto (reduced-precision sweeps are sample-checked on the device and redone when they fail), 3 parity,
    2 / 1 reduced.  Returns (sums[3] float64 device tensor {zz, pz, pp}, grad_unscaled [rows, D])."""
    L.require(z_all, "z", dims=(2,)); L.require(prior_all, "prior_z", dims=(2,))
    if z_all.shape != prior_all.shape:
        raise RuntimeError("vaekit: z %s and prior_z %s differ in shape" % (tuple(z_all.shape), tuple(prior_all.shape)))
    z, p = z_all.contiguous(), prior_all.contiguous()
    N, D = z.shape
    dev = z.device
    need = L.ws_bytes("mmd", N, D)
    if need < 0:
        raise RuntimeError("vaekit: MMD shape N=%d D=%d is not supported" % (N, D))
    ws = L.workspace(dev, need)
    sums = torch.empty(3, dtype=torch.float64, device=dev)
    grad = torch.empty((row_end - row_begin, D), dtype=torch.float32, device=dev)
    diag = torch.empty(4, dtype=torch.float32, device=dev) if MMD_DIAG is not None else None
    with L.on(dev):
        L.call("mmd_fwd", L.ptr(z), L.ptr(p), N, D, kernel_type, z_var, passes, row_begin, row_end, L.ptr(sums),
               L.ptr(grad), L.ptr(diag), L.ptr(ws), ws.numel(), L.stream_ptr())
    if diag is not None:
        MMD_DIAG.append(diag)
    return sums, grad


def mmd_combine(sums, N, kernel_type, reg_weight):
    out = torch.empty(4, dtype=torch.float32, device=sums.device)
    with L.on(sums.device):
        L.call("mmd_combine", L.ptr(sums), N, kernel_type, reg_weight, L.ptr(out), L.stream_ptr())
    return out


def scale_by_scalar(t, g, factor):
    out = torch.empty_like(t)
    with L.on(t.device):
        L.call("scale_by_scalar", L.ptr(t), L.ptr(g), factor, t.numel(), L.ptr(out), L.stream_ptr())
    return out


class _Mmd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, prior, kernel_type, z_var, reg_weight, passes):
        N = z.shape[0]
        sums, grad = mmd_sweep(z, prior, kernel_type, z_var, 0, N, passes)
        out = mmd_combine(sums, N, kernel_type, reg_weight)
        ctx.save_for_backward(grad)
        ctx.factor = 4.0 * reg_weight * (1.0 / (float(N) * float(N)) if kernel_type == 0 else 1.0)
        ctx.set_materialize_grads(False)
        terms = out[1:].detach()
        ctx.mark_non_differentiable(terms)
        return out[0], terms

    @staticmethod
    def backward(ctx, g_mmd, _g_terms):
        (grad,) = ctx.saved_tensors
        if g_mmd is None:
            return None, None, None, None, None, None
        return scale_by_scalar(grad, g_mmd.contiguous(), ctx.factor), None, None, None, None, None


def mmd(z, prior_z, reg_weight, kernel_type, z_var, passes=0):
    """wae_mmd.py:193-204 with the prior sample given: returns (mmd, terms[3] = {pp, zz, pz} kernel means/sums).
    Value and d mmd/d z come out of one tcgen05 Gram sweep; backward only rescales the saved gradient."""
    if kernel_type not in KERNEL_TYPES:
        raise ValueError('Undefined kernel type.')
    return _Mmd.apply(z, prior_z, KERNEL_TYPES[kernel_type], float(z_var), float(reg_weight), int(passes))


# ---------------------------------------------------------------------------
# K4b: the kernel matrix on its own (compute_kernel / compute_rbf / compute_inv_mult_quad called directly)
# ---------------------------------------------------------------------------
def _pair_fwd(x1, s1, x2, s2, N1, N2, D, kt, z_var, eps):
    dev = x1.device
    kmat = torch.empty((N1, N2), dtype=torch.float32, device=dev)
    imq = torch.empty(1, dtype=torch.float32, device=dev) if kt == 1 else None
    ws = L.workspace(dev)
    with L.on(dev):
        L.call("pair_kernel_fwd", L.ptr(x1), L.i64x3(s1), L.ptr(x2), L.i64x3(s2), N1, N2, D, kt, z_var, eps, L.ptr(kmat),
               L.ptr(imq), L.ptr(ws), ws.numel(), L.stream_ptr())
    return kmat, imq


def _upstream(g, kt):
    """(tensor, row stride, column stride) of the upstream gradient: [N1, N2] for rbf, a scalar for imq."""
    g = g.contiguous()
    return (g, 0, 0) if kt == 1 else (g, g.stride(0), g.stride(1))


class _PairKernelRows(torch.autograd.Function):
    """k(x1_i, x2_j) for x1 [N1, D], x2 [N2, D]: what compute_kernel computes after its unsqueeze/expand."""

    @staticmethod
    def forward(ctx, x1, x2, kt, z_var, eps):
        N1, D = x1.shape
        N2 = x2.shape[0]
        kmat, imq = _pair_fwd(x1, (x1.stride(0), 0, x1.stride(1)), x2, (0, x2.stride(0), x2.stride(1)), N1, N2, D, kt,
                              z_var, eps)
        ctx.save_for_backward(x1, x2, kmat)
        ctx.cfg = (kt, z_var, eps)
        return kmat if kt == 0 else imq[0]

    @staticmethod
    def backward(ctx, g):
        x1, x2, kmat = ctx.saved_tensors
        kt, z_var, eps = ctx.cfg
        N1, D = x1.shape
        N2 = x2.shape[0]
        g, g_sr, g_sc = _upstream(g, kt)
        g1 = torch.empty((N1, D), dtype=torch.float32, device=x1.device) if ctx.needs_input_grad[0] else None
        g2 = torch.empty((N2, D), dtype=torch.float32, device=x1.device) if ctx.needs_input_grad[1] else None
        with L.on(x1.device):
            L.call("pair_kernel_bwd_rows", L.ptr(x1), x1.stride(0), x1.stride(1), L.ptr(x2), x2.stride(0), x2.stride(1), N1,
                   N2, D, kt, z_var, eps, L.ptr(kmat), L.ptr(g), g_sr, g_sc, L.ptr(g1), L.ptr(g2), L.stream_ptr())
        return g1, g2, None, None, None


class _PairKernelDense(torch.autograd.Function):
    """The same kernels on general [N1, N2, D] operands (any strides)."""

    @staticmethod
    def forward(ctx, x1, x2, kt, z_var, eps):
        N1, N2, D = x1.shape
        kmat, imq = _pair_fwd(x1, x1.stride(), x2, x2.stride(), N1, N2, D, kt, z_var, eps)
        ctx.save_for_backward(x1, x2, kmat)
        ctx.cfg = (kt, z_var, eps)
        return kmat if kt == 0 else imq[0]

    @staticmethod
    def backward(ctx, g):
        x1, x2, kmat = ctx.saved_tensors
        kt, z_var, eps = ctx.cfg
        N1, N2, D = x1.shape
        g, g_sr, g_sc = _upstream(g, kt)
        g1 = torch.empty((N1, N2, D), dtype=torch.float32, device=x1.device) if ctx.needs_input_grad[0] else None
        g2 = torch.empty((N1, N2, D), dtype=torch.float32, device=x1.device) if ctx.needs_input_grad[1] else None
        with L.on(x1.device):
            L.call("pair_kernel_bwd_dense", L.ptr(x1), L.i64x3(x1.stride()), L.ptr(x2), L.i64x3(x2.stride()), N1, N2, D,
                   kt, z_var, eps, L.ptr(kmat), L.ptr(g), g_sr, g_sc, L.ptr(g1), L.ptr(g2), L.stream_ptr())
        return g1, g2, None, None, None


def pair_kernel(x1, x2, kernel_type, z_var, eps=1e-7):
    """compute_rbf / compute_inv_mult_quad (wae_mmd.py:153-191) on the reference's [N1, N2, D] operands.  When they are
    the expand() views compute_kernel builds (x1 constant along axis 1, x2 along axis 0) the work is done on the
    [N, D] bases and the gradient comes back already summed over the broadcast axis; otherwise the dense form runs.
    Returns [N1, N2] for 'rbf' and the 0-dim `sum - diag sum` for 'imq', like the reference."""
    if kernel_type not in KERNEL_TYPES:
        raise ValueError('Undefined kernel type.')
    kt = KERNEL_TYPES[kernel_type]
    L.require(x1, "x1", dims=(3,)); L.require(x2, "x2", dims=(3,))
    if x1.shape != x2.shape:
        x1, x2 = torch.broadcast_tensors(x1, x2)
    if 0 in x1.shape:
        raise RuntimeError("vaekit: empty operand for the kernel matrix")
    if x1.shape[1] > 1 and x1.shape[0] > 1 and x1.stride(1) == 0 and x2.stride(0) == 0:
        return _PairKernelRows.apply(x1[:, 0, :], x2[0, :, :], kt, float(z_var), float(eps))
    return _PairKernelDense.apply(x1, x2, kt, float(z_var), float(eps))


# ---------------------------------------------------------------------------
# K6
# ---------------------------------------------------------------------------
def swd_sweep(z, prior, proj, p, want_grad=True):
    """Projection GEMM + segmented radix sorts + loss (+ d/dz) for the projections in `proj` [S, D] (vk_swd_fwd).
    Returns (sum[1] float64 device tensor = sum (sorted diff)^p, grad_unscaled [N, D] or None)."""
    L.require(z, "z", dims=(2,)); L.require(prior, "prior_z", dims=(2,)); L.require(proj, "projections", dims=(2,))
    if z.shape != prior.shape or proj.shape[1] != z.shape[1]:
        raise RuntimeError("vaekit: z %s, prior_z %s and projections %s do not fit together" %
                           (tuple(z.shape), tuple(prior.shape), tuple(proj.shape)))
    zc, pc, th = z.contiguous(), prior.contiguous(), proj.contiguous()
    N, D = zc.shape
    S = th.shape[0]
    if N * S == 0:
        raise RuntimeError("vaekit: empty input to the sliced-Wasserstein distance")
    dev = zc.device
    ws = L.workspace(dev, L.ws_bytes("swd", N, S))
    total = torch.empty(1, dtype=torch.float64, device=dev)
    grad = torch.empty((N, D), dtype=torch.float32, device=dev) if want_grad else None
    with L.on(dev):
        L.call("swd_fwd", L.ptr(zc), L.ptr(pc), L.ptr(th), N, D, S, p, L.ptr(total), L.ptr(grad), L.ptr(ws), ws.numel(),
               L.stream_ptr())
    return total, grad


def mean_scale(total, count, scale):
    out = torch.empty(1, dtype=torch.float32, device=total.device)
    with L.on(total.device):
        L.call("mean_scale", L.ptr(total), float(count), scale, L.ptr(out), L.stream_ptr())
    return out[0]


class _Swd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, prior, proj, p, reg_weight):
        total, grad = swd_sweep(z, prior, proj, p, want_grad=ctx.needs_input_grad[0])
        count = float(z.shape[0]) * float(proj.shape[0])
        if grad is not None:
            ctx.save_for_backward(grad)
        ctx.factor = reg_weight / count
        ctx.set_materialize_grads(False)
        return mean_scale(total, count, reg_weight)

    @staticmethod
    def backward(ctx, g):
        if g is None or not ctx.saved_tensors:
            return None, None, None, None, None
        return scale_by_scalar(ctx.saved_tensors[0], g.contiguous(), ctx.factor), None, None, None, None


def swd(z, prior_z, proj, p, reg_weight):
    """swae.py:165-179 with the prior sample and the [S, D] projection directions given."""
    return _Swd.apply(z, prior_z, proj, float(p), float(reg_weight))


def swd_project(x, proj, out=None):
    """Projection-major projections out[s, n] = x[n, :] . proj[s, :] (the transposed z.matmul(proj) of swae.py:171-176)
    with the fp32-grade tcgen05 GEMM (vk_gemm_nt, transposed store).  `out` may be a [S, n] view with stride(1) == 1
    (e.g. one channel of an exchange buffer)."""
    L.require(x, "x", dims=(2,)); L.require(proj, "projections", dims=(2,))
    if x.shape[1] != proj.shape[1]:
        raise RuntimeError("vaekit: rows %s and projections %s do not fit together" % (tuple(x.shape), tuple(proj.shape)))
    n, D = x.shape
    S = proj.shape[0]
    if out is None:
        out = torch.empty((S, n), dtype=torch.float32, device=x.device)
    if tuple(out.shape) != (S, n) or (n > 1 and out.stride(1) != 1) or out.dtype != torch.float32:
        raise RuntimeError("vaekit: swd_project output must be a [S, n] fp32 view contiguous along n")
    if n * S == 0:
        return out
    ws = L.workspace(x.device, L.ws_bytes("gemm_nt"))
    with L.on(x.device):
        L.call("gemm_nt", L.ptr(x), x.stride(0), x.stride(1), L.ptr(proj), proj.stride(0), proj.stride(1), L.ptr(out), 1,
               out.stride(0), n, S, D, L.ptr(ws), ws.numel(), L.stream_ptr())
    return out


def swd_segments(lat, pri, p, want_grad=True):
    """Sorts + loss (+ d/dlat through the permutation) of [S, N] projection-major segments (vk_swd_segments_fwd).
    Both inputs are consumed: on return `lat` holds d sum / d lat in sample order (when want_grad).
    Returns sum[1] float64 device tensor."""
    L.require(lat, "lat", dims=(2,)); L.require(pri, "pri", dims=(2,))
    if lat.shape != pri.shape or not lat.is_contiguous() or not pri.is_contiguous():
        raise RuntimeError("vaekit: swd_segments needs two contiguous [S, N] tensors of the same shape")
    S, N = lat.shape
    total = torch.zeros(1, dtype=torch.float64, device=lat.device)
    if S * N == 0:
        return total
    ws = L.workspace(lat.device, L.ws_bytes("swd_segments", N, S))
    with L.on(lat.device):
        L.call("swd_segments_fwd", L.ptr(lat), L.ptr(pri), N, S, p, L.ptr(total), 1 if want_grad else 0, L.ptr(ws),
               ws.numel(), L.stream_ptr())
    return total


def swd_project_back(g_lat, proj):
    """dz[n, :] = sum_s g_lat[s, n] * proj[s, :]  (the gradient GEMM of swae.py:171 through vk_gemm_nt)."""
    L.require(g_lat, "g_lat", dims=(2,)); L.require(proj, "projections", dims=(2,))
    S, n = g_lat.shape
    D = proj.shape[1]
    out = torch.empty((n, D), dtype=torch.float32, device=g_lat.device)
    if n == 0:
        return out
    if S == 0:
        return out.zero_()
    ws = L.workspace(g_lat.device, L.ws_bytes("gemm_nt"))
    with L.on(g_lat.device):
        L.call("gemm_nt", L.ptr(g_lat), g_lat.stride(1), g_lat.stride(0), L.ptr(proj), proj.stride(1), proj.stride(0),
               L.ptr(out), D, 1, n, D, S, L.ptr(ws), ws.numel(), L.stream_ptr())
    return out


def swd_project_cols(x, proj, c_cols):
    """swd_project with every output column stored at its own base address: c_cols is an int64 CUDA tensor [S] of
    device pointers, c_cols[s] = address of out(s, 0) (contiguous along the rows of x).  The addresses may be
    peer-mapped memory of other GPUs (vk_gemm_nt_tab)."""
    L.require(x, "x", dims=(2,)); L.require(proj, "projections", dims=(2,))
    n, D = x.shape
    S = proj.shape[0]
    if c_cols.dtype != torch.int64 or not c_cols.is_cuda or c_cols.numel() != S:
        raise RuntimeError("vaekit: column table must be an int64 CUDA tensor with one address per projection")
    if n * S == 0:
        return
    ws = L.workspace(x.device, L.ws_bytes("gemm_nt"))
    with L.on(x.device):
        L.call("gemm_nt_tab", None, L.ptr(x), x.stride(0), x.stride(1), L.ptr(proj), proj.stride(0), proj.stride(1),
               L.ptr(c_cols), None, 1, 0, n, S, D, L.ptr(ws), ws.numel(), L.stream_ptr())


def swd_project_back_cols(a_cols, n, proj):
    """swd_project_back with every d/dlat column read from its own base address (a_cols[s] = address of g_lat(s, 0),
    n contiguous floats, possibly in a peer GPU's memory)."""
    L.require(proj, "projections", dims=(2,))
    S, D = proj.shape
    if a_cols.dtype != torch.int64 or not a_cols.is_cuda or a_cols.numel() != S:
        raise RuntimeError("vaekit: column table must be an int64 CUDA tensor with one address per projection")
    out = torch.empty((n, D), dtype=torch.float32, device=proj.device)
    if n == 0:
        return out
    ws = L.workspace(proj.device, L.ws_bytes("gemm_nt"))
    with L.on(proj.device):
        L.call("gemm_nt_tab", L.ptr(a_cols), None, 1, 0, L.ptr(proj), proj.stride(1), proj.stride(0), None, L.ptr(out), D, 1,
               n, D, S, L.ptr(ws), ws.numel(), L.stream_ptr())
    return out


def seg_sort(keys, want_perm=False):
    """Ascending sort of every row of a [segments, n] fp32 CUDA tensor with the library's radix sort.
    Returns (sorted, perm or None); perm[s, i] (int64) is the source column of sorted[s, i]."""
    L.require(keys, "keys", dims=(2,))
    out = keys.contiguous().clone()
    S, n = out.shape
    if S * n == 0:
        return out, (torch.empty((S, n), dtype=torch.int64, device=out.device) if want_perm else None)
    dev = out.device
    perm = torch.empty((S, n), dtype=torch.int32, device=dev) if want_perm else None
    ws = L.workspace(dev, L.ws_bytes("seg_sort", S, n))
    with L.on(dev):
        L.call("seg_sort", L.ptr(out), L.ptr(perm), S, n, L.ptr(ws), ws.numel(), L.stream_ptr())
    return out, (perm.long() if want_perm else None)


def gemm_nt(a, b):
    """a [N, K] @ b[J, K]^T -> [N, J] with the library's fp32-grade tcgen05 GEMM (any strides; J, K small)."""
    L.require(a, "a", dims=(2,))
    L.require(b, "b", dims=(2,))
    if a.shape[1] != b.shape[1]:
        raise ValueError("gemm_nt: inner dimensions differ: %s vs %s" % (tuple(a.shape), tuple(b.shape)))
    N, K = a.shape
    J = b.shape[0]
    out = torch.empty((N, J), dtype=torch.float32, device=a.device)
    if N * J == 0:
        return out
    ws = L.workspace(a.device, L.ws_bytes("gemm_nt"))
    with L.on(a.device):
        L.call("gemm_nt", L.ptr(a), a.stride(0), a.stride(1), L.ptr(b), b.stride(0), b.stride(1), L.ptr(out), J, 1,
               N, J, K, L.ptr(ws), ws.numel(), L.stream_ptr())
    return out


# ---------------------------------------------------------------------------
# K5
# ---------------------------------------------------------------------------
class _BetaTc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, mu, log_var, M_N, alpha, beta, gamma_anneal):
        for t, n in ((z, "z"), (mu, "mu"), (log_var, "log_var")):
            L.require(t, n, dims=(2,))
        if not (z.shape == mu.shape == log_var.shape):
            raise RuntimeError("vaekit: z %s, mu %s and log_var %s differ in shape" %
                               (tuple(z.shape), tuple(mu.shape), tuple(log_var.shape)))
        B, D = z.shape
        if B < 2:
            raise RuntimeError("vaekit: the Beta-TC decomposition needs a batch of at least 2")
        zc, mc, lc = z.contiguous(), mu.contiguous(), log_var.contiguous()
        dev = z.device
        lqz = torch.empty(2 * B, dtype=torch.float32, device=dev)     # value, then its fp32 residual
        lse = torch.empty((B, D), dtype=torch.float32, device=dev)
        out = torch.empty(4, dtype=torch.float32, device=dev)
        ws = L.workspace(dev, L.ws_bytes("tc", B, D))
        with L.on(dev):
            L.call("tc_fwd", L.ptr(zc), L.ptr(mc), L.ptr(lc), B, D, M_N, alpha, beta, gamma_anneal, L.ptr(lqz), L.ptr(lse),
                   L.ptr(out), L.ptr(ws), ws.numel(), L.stream_ptr())
        ctx.save_for_backward(zc, mc, lc, lqz, lse)
        ctx.cfg = (B, D, M_N, alpha, beta, gamma_anneal)
        ctx.set_materialize_grads(False)
        terms = out[1:].detach()
        ctx.mark_non_differentiable(terms)
        return out[0], terms

    @staticmethod
    def backward(ctx, g, _g_terms):
        if g is None:
            return (None,) * 7
        zc, mc, lc, lqz, lse = ctx.saved_tensors
        B, D, M_N, alpha, beta, gamma_anneal = ctx.cfg
        dev = zc.device
        g_z, g_mu, g_lv = torch.empty_like(zc), torch.empty_like(mc), torch.empty_like(lc)
        ws = L.workspace(dev, L.ws_bytes("tc", B, D))
        with L.on(dev):
            L.call("tc_bwd", L.ptr(zc), L.ptr(mc), L.ptr(lc), L.ptr(lqz), L.ptr(lse), B, D, M_N, alpha, beta, gamma_anneal,
                   L.ptr(g.contiguous()), L.ptr(g_z), L.ptr(g_mu), L.ptr(g_lv), L.ptr(ws), ws.numel(), L.stream_ptr())
        return g_z, g_mu, g_lv, None, None, None, None


def betatc_rows_fwd(z_all, mu_all, lv_all, row_begin, row_end, M_N, alpha, beta, gamma_anneal):
    """Forward row sweep of anchor rows [row_begin, row_end) against all B columns (vk_tc_fwd_rows).  Returns
    (out[4] = the range's share of {latent loss, kld, tc, mi}, lqz [2B], lse [B, D]) with only the range's rows valid."""
    for t, n in ((z_all, "z"), (mu_all, "mu"), (lv_all, "log_var")):
        L.require(t, n, dims=(2,))
    B, D = z_all.shape
    dev = z_all.device
    zc, mc, lc = z_all.contiguous(), mu_all.contiguous(), lv_all.contiguous()
    lqz = torch.zeros(2 * B, dtype=torch.float32, device=dev)
    lse = torch.zeros((B, D), dtype=torch.float32, device=dev)
    out = torch.empty(4, dtype=torch.float32, device=dev)
    ws = L.workspace(dev, L.ws_bytes("tc", B, D))
    with L.on(dev):
        L.call("tc_fwd_rows", L.ptr(zc), L.ptr(mc), L.ptr(lc), B, D, int(row_begin), int(row_end), M_N, alpha, beta,
               gamma_anneal, L.ptr(lqz), L.ptr(lse), L.ptr(out), L.ptr(ws), ws.numel(), L.stream_ptr())
    return out, lqz, lse


def betatc_rows_bwd(z_all, mu_all, lv_all, lqz_all, lse_all, row_begin, row_end, M_N, alpha, beta, gamma_anneal, g):
    """Backward of the rows [row_begin, row_end): (g_z, g_mu, g_lv) [rows, D] (vk_tc_bwd_rows); lqz_all / lse_all must
    hold every row's forward values."""
    B, D = z_all.shape
    dev = z_all.device
    g_z, g_mu, g_lv = (torch.empty((B, D), dtype=torch.float32, device=dev) for _ in range(3))
    ws = L.workspace(dev, L.ws_bytes("tc", B, D))
    with L.on(dev):
        L.call("tc_bwd_rows", L.ptr(z_all), L.ptr(mu_all), L.ptr(lv_all), L.ptr(lqz_all), L.ptr(lse_all), B, D,
               int(row_begin), int(row_end), M_N, alpha, beta, gamma_anneal, L.ptr(g.contiguous()), L.ptr(g_z), L.ptr(g_mu),
               L.ptr(g_lv), L.ptr(ws), ws.numel(), L.stream_ptr())
    return g_z[row_begin:row_end], g_mu[row_begin:row_end], g_lv[row_begin:row_end]


def betatc_terms(z, mu, log_var, M_N, alpha, beta, gamma_anneal):
    """betatc_vae.py:165-205 without the reconstruction term: returns (alpha*mi + beta*tc + gamma_anneal*kld,
    terms[3] = {kld, tc, mi})."""
    return _BetaTc.apply(z, mu, log_var, float(M_N), float(alpha), float(beta), float(gamma_anneal))


# ---------------------------------------------------------------------------
# K8: Gumbel-softmax + categorical KL
# ---------------------------------------------------------------------------
class _GumbelSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, temp, eps, seed, offset, u_in, want_u):
        L.require(q, "logits")
        if q.dim() < 1 or q.numel() == 0:
            raise RuntimeError("vaekit: empty logits")
        qc = q.contiguous()
        C = qc.shape[-1]
        R = qc.numel() // C
        s = torch.empty_like(qc)
        u_out = torch.empty_like(qc) if want_u else None
        uc = None if u_in is None else L.require(u_in, "uniforms").contiguous()
        with L.on(qc.device):
            L.call("gumbel_softmax_fwd", L.ptr(qc), R, C, temp, eps, seed, offset, L.ptr(uc), L.ptr(s), L.ptr(u_out),
                   L.stream_ptr())
        ctx.save_for_backward(s)
        ctx.cfg = (R, C, temp)
        if u_out is not None:
            ctx.mark_non_differentiable(u_out)
        return s, u_out

    @staticmethod
    def backward(ctx, g_s, _g_u):
        if g_s is None:
            return (None,) * 7
        s, = ctx.saved_tensors
        R, C, temp = ctx.cfg
        g_q = torch.empty_like(s)
        with L.on(s.device):
            L.call("gumbel_softmax_bwd", L.ptr(s), L.ptr(g_s.contiguous()), R, C, temp, L.ptr(g_q), L.stream_ptr())
        return g_q, None, None, None, None, None, None


def gumbel_softmax(q, temp, eps=1e-7, *, rng=None, uniforms=None, want_uniforms=False):
    """cat_vae.py:118-131 on logits [..., C]: softmax((q + gumbel) / temp, -1), same shape as q.  The uniforms come from the
    contract Philox stream (torch.manual_seed governs them) unless `uniforms` is given.  Returns (s, uniforms or None)."""
    if rng is None:
        if q.is_cuda and uniforms is None and _DEVICE_KEYS.get(q.device.index) is not None:
            raise RuntimeError("vaekit: gumbel_softmax has no device-key entry point yet; pass uniforms= inside device_rng()")
        rng = next_philox(q.device, q.numel()) if q.is_cuda and uniforms is None else (0, 0)
    return _GumbelSoftmax.apply(q, float(temp), float(eps), rng[0], rng[1], uniforms, bool(want_uniforms))


class _CatKlRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, eps):
        L.require(q, "logits", dims=(3,))
        if q.numel() == 0:
            raise RuntimeError("vaekit: empty logits")
        qc = q.contiguous()
        B, Ld, C = qc.shape
        rows = torch.empty(B, dtype=torch.float32, device=qc.device)
        with L.on(qc.device):
            L.call("cat_kl_rows_fwd", L.ptr(qc), B, Ld, C, eps, L.ptr(rows), L.stream_ptr())
        ctx.save_for_backward(qc)
        ctx.eps = eps
        return rows

    @staticmethod
    def backward(ctx, g_rows):
        qc, = ctx.saved_tensors
        B, Ld, C = qc.shape
        g_q = torch.empty_like(qc)
        with L.on(qc.device):
            L.call("cat_kl_rows_bwd", L.ptr(qc), L.ptr(g_rows.contiguous()), B, Ld, C, ctx.eps, L.ptr(g_q), L.stream_ptr())
        return g_q, None


def cat_kl_rows(q, eps=1e-7):
    """Per-sample KL(softmax(q) || uniform) of cat_vae.py:156-169 for logits [B, latent_dim, C]: [B] rows whose batch mean
    is the reference's kld_loss."""
    return _CatKlRows.apply(q, float(eps))


# ---------------------------------------------------------------------------
# K9: VampPrior term
# ---------------------------------------------------------------------------
class _GaussLogqRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, mu, log_var):
        for t, n in ((z, "z"), (mu, "mu"), (log_var, "log_var")):
            L.require(t, n, dims=(2,))
        if z.shape != mu.shape or z.shape != log_var.shape or z.numel() == 0:
            raise RuntimeError("vaekit: z / mu / log_var must share a non-empty [B, D] shape")
        zc, mc, vc = z.contiguous(), mu.contiguous(), log_var.contiguous()
        B, D = zc.shape
        rows = torch.empty(B, dtype=torch.float32, device=zc.device)
        with L.on(zc.device):
            L.call("gauss_logq_rows_fwd", L.ptr(zc), L.ptr(mc), L.ptr(vc), B, D, L.ptr(rows), L.stream_ptr())
        ctx.save_for_backward(zc, mc, vc)
        return rows

    @staticmethod
    def backward(ctx, g_rows):
        zc, mc, vc = ctx.saved_tensors
        B, D = zc.shape
        g_z, g_mu, g_lv = torch.empty_like(zc), torch.empty_like(zc), torch.empty_like(zc)
        with L.on(zc.device):
            L.call("gauss_logq_rows_bwd", L.ptr(zc), L.ptr(mc), L.ptr(vc), L.ptr(g_rows.contiguous()), B, D, L.ptr(g_z),
                   L.ptr(g_mu), L.ptr(g_lv), L.stream_ptr())
        return g_z, g_mu, g_lv


def gauss_logq_rows(z, mu, log_var):
    """vampvae.py:136-138 per sample: sum_d -0.5 (log_var + (z - mu)^2) / exp(log_var), [B]; its batch mean is E_log_q_z."""
    return _GaussLogqRows.apply(z, mu, log_var)


class _VampPrior(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, prior_mu, prior_log_var):
        L.require(z, "z", dims=(2,))
        L.require(prior_mu, "prior_mu", dims=(2,))
        L.require(prior_log_var, "prior_log_var", dims=(2,))
        if prior_mu.shape != prior_log_var.shape or prior_mu.shape[1] != z.shape[1] or z.numel() == 0 or prior_mu.numel() == 0:
            raise RuntimeError("vaekit: vamp prior needs z [B, D] and prior_mu / prior_log_var [K, D]")
        zc, mc, vc = z.contiguous(), prior_mu.contiguous(), prior_log_var.contiguous()
        B, D = zc.shape
        K = mc.shape[0]
        t = torch.empty(B, K, dtype=torch.float32, device=zc.device)
        lse = torch.empty(B, dtype=torch.float32, device=zc.device)
        with L.on(zc.device):
            L.call("vamp_prior_fwd", L.ptr(zc), L.ptr(mc), L.ptr(vc), B, K, D, L.ptr(t), L.ptr(lse), L.stream_ptr())
        ctx.save_for_backward(zc, mc, vc, t, lse)
        return lse

    @staticmethod
    def backward(ctx, g_rows):
        zc, mc, vc, t, lse = ctx.saved_tensors
        B, D = zc.shape
        K = mc.shape[0]
        g_z, g_m, g_v = torch.empty_like(zc), torch.empty_like(mc), torch.empty_like(vc)
        with L.on(zc.device):
            L.call("vamp_prior_bwd", L.ptr(zc), L.ptr(mc), L.ptr(vc), L.ptr(t), L.ptr(lse), L.ptr(g_rows.contiguous()), B, K, D,
                   L.ptr(g_z), L.ptr(g_m), L.ptr(g_v), L.stream_ptr())
        return g_z, g_m, g_v


def vamp_prior_rows(z, prior_mu, prior_log_var):
    """vampvae.py:149-158 per sample: logsumexp over the K pseudo-input components of the (reference's un-normalised)
    Gaussian log density minus log K, [B]; its batch mean is E_log_p_z.  The [B, K, D] broadcast is not materialised."""
    return _VampPrior.apply(z, prior_mu, prior_log_var)
