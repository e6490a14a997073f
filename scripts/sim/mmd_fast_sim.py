"""CPU emulation of the operand roundings of the MMD 'fast' mode (no GPU needed): centred data, 1-pass bf16 distance GEMM,
centred coefficients c' = c - c0 in 1-pass bf16 for the gradient GEMM.  fp32 accumulation is emulated with torch fp32
matmuls (round-to-nearest, not the tensor core's truncation: that part is only measurable on the B200)."""
import sys
import torch

torch.set_num_threads(8)


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def split(x):
    h = bf(x); return h, bf(x - h)


def ref64(z, p, kt, z_var):
    z = z.double().requires_grad_(True); p = p.double()
    N, D = z.shape

    def K(a, b):
        d2 = (a * a).sum(1)[:, None] + (b * b).sum(1)[None] - 2 * a @ b.t()
        if kt == "rbf":
            return torch.exp(-d2 / (2 * D * D * z_var))
        C = 2 * D * z_var
        return C / (1e-7 + C + d2)
    if kt == "rbf":
        zz, pz, pp = K(z, z).mean(), K(p, z).mean(), K(p, p).mean()
    else:
        kzz, kpz, kpp = K(z, z), K(p, z), K(p, p)
        zz, pz, pp = kzz.sum() - kzz.diag().sum(), kpz.sum() - kpz.diag().sum(), kpp.sum() - kpp.diag().sum()
    mmd = pp + zz - 2 * pz
    g, = torch.autograd.grad(mmd, z)
    return g, (zz.item(), pz.item(), pp.item())


def fast(z, p, kt, z_var, g1_passes, g2_passes, center_data=True, center_c=True):
    N, D = z.shape
    m = (z.mean(0) + p.mean(0)) * 0.5 if center_data else torch.zeros(D)
    zc, pc = z - m, p - m
    a = 1.0 / (2 * D * D * z_var); C = 2 * D * z_var

    def dot(x, y, passes):
        xh, xl = split(x); yh, yl = split(y)
        s = xh @ yh.t()
        if passes == 3:
            s = (xl @ yh.t() + xh @ yl.t()) + s
        return s

    def kern(d2):
        if kt == "rbf":
            k = torch.exp(-d2 * a); return k, k
        k = 1.0 / (C + 1e-7 + d2); return k * C, k * k          # value, coefficient (k/C)^2

    nz, npn = (zc * zc).sum(1), (pc * pc).sum(1)
    dbar_zz = 2 * nz.mean(); dbar_pz = nz.mean() + npn.mean()
    c0zz = kern(dbar_zz)[1] if center_c else torch.tensor(0.0)
    c0pz = kern(dbar_pz)[1] if center_c else torch.tensor(0.0)
    kzz, czz = kern(nz[:, None] + nz[None] - 2 * dot(zc, zc, g1_passes))
    kzp, czp = kern(nz[:, None] + npn[None] - 2 * dot(zc, pc, g1_passes))     # rows z, cols prior

    def gemm2(c, y, passes):
        ch, cl = split(c); yh, yl = split(y)
        o = ch @ yh
        if passes == 3:
            o = (cl @ yh + ch @ yl) + o
        return o
    if kt == "imq":
        czz = czz.clone(); czz.fill_diagonal_(0); czp = czp.clone(); czp.fill_diagonal_(0)
    czz_c, czp_c = czz - c0zz, czp - c0pz
    if kt == "imq" and center_c:          # the diagonal carries no coefficient at all: c' = -c0 there would be wrong
        czz_c.fill_diagonal_(0); czp_c.fill_diagonal_(0)
    O = gemm2(czz_c, zc, g2_passes) - gemm2(czp_c, pc, g2_passes)
    rs = czz_c.sum(1) - czp_c.sum(1)
    nvalid = N - (1 if (kt == "imq" and center_c) else 0)
    # analytic part: sum_j c0 (z_i - y_j) over the columns that carry a coefficient
    Sz, Sp = zc.double().sum(0), pc.double().sum(0)
    if kt == "imq" and center_c:
        ana = (c0zz.double() * (nvalid * zc.double() - (Sz[None] - zc.double()))
               - c0pz.double() * (nvalid * zc.double() - (Sp[None] - pc.double())))
    else:
        ana = c0zz.double() * (N * zc.double() - Sz[None]) - c0pz.double() * (N * zc.double() - Sp[None])
    gsum = (zc * rs[:, None] - O).double() + ana
    scale = (-2 * a * 2 / (N * N)) if kt == "rbf" else (-2 * C * 2)
    g = gsum * scale
    if kt == "rbf":
        sums = (kzz.double().mean().item(), kzp.double().mean().item())
    else:
        sums = ((kzz.double().sum() - kzz.double().diag().sum()).item(), (kzp.double().sum() - kzp.double().diag().sum()).item())
    return g, sums


if __name__ == "__main__":
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    for D, shift, sig in ((128, 0.25, 1.0), (128, 0.0, 1.0), (64, 0.0, 1.0), (16, 0.5, 1.0), (10, 0.5, 1.0), (2, 0.5, 1.0), (128, 0.25, 3.0)):
        g0 = torch.Generator().manual_seed(1)
        z = sig * torch.randn(N, D, generator=g0) + shift; p = torch.randn(N, D, generator=g0)
        for kt in ("rbf", "imq"):
            gref, sref = ref64(z, p, kt, 2.0)
            den = gref.abs().max()
            out = []
            for (p1, p2, cd, cc) in ((3, 3, False, False), (3, 3, True, True), (3, 1, True, True), (1, 1, True, True), (1, 3, True, True), (1, 1, False, True), (1, 1, True, False)):
                g, s = fast(z, p, kt, 2.0, p1, p2, cd, cc)
                e = ((g - gref).abs().max() / den).item()
                es = max(abs(s[0] - sref[0]) / abs(sref[0]), abs(s[1] - sref[1]) / abs(sref[1]))
                out.append("g1=%d g2=%d cd=%d cc=%d: grad %.1e sums %.1e" % (p1, p2, cd, cc, e, es))
            print("N=%d D=%d shift=%.2f sig=%.1f %s\n   " % (N, D, shift, sig, kt) + "\n   ".join(out), flush=True)
