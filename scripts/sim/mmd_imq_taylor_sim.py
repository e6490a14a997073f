"""CPU emulation: could the IMQ sweep run in a reduced-precision mode if the coefficient's first-order Taylor term in d^2 were
handled analytically (like c0 is now)?

  c(d2) = 1 / (C + eps + d2)^2 ;  c = c0 + c1 (d2 - dbar) + c''   with c0 = c(dbar), c1 = c'(dbar)
  tensor path:  c'' from the ROUNDED d2 (bf16 operands), rounded to bf16 (hi [+ lo]) for the gradient GEMM
  analytic:     sum_j [c0 + c1 (d2_ij - dbar)] (z_i - y_j) in float64 from moments (row effect, column effect and the bilinear
                effect -2 c1 x_i.y_j need S_y, sum_j |y_j|^2 y_j and M = sum_j y_j y_j^T); here evaluated directly in float64
fp32 accumulation is emulated with torch fp32 matmuls (round to nearest; the tensor core truncates: only measurable on the B200).
argv: N"""
import sys
import torch

torch.set_num_threads(8)
bf = lambda x: x.to(torch.bfloat16).to(torch.float32)


def split(x):
    h = bf(x); return h, bf(x - h)


def ref64(z, p, z_var):
    z = z.double().requires_grad_(True); p = p.double()
    N, D = z.shape
    C = 2 * D * z_var

    def K(a, b):
        d2 = (a * a).sum(1)[:, None] + (b * b).sum(1)[None] - 2 * a @ b.t()
        return C / (1e-7 + C + d2)
    kzz, kpz = K(z, z), K(p, z)
    val = (kzz.sum() - kzz.diag().sum()) - 2 * (kpz.sum() - kpz.diag().sum())
    g, = torch.autograd.grad(val, z)
    return g


def run(z, p, z_var, g1, g2, order):
    """g1: 1 plain bf16, 2 row operand hi+lo, 3 both hi+lo; g2: 1 coefficient hi x column hi, 3 full; order: Taylor terms
    removed from the tensor path (0: c0 only, 1: c0 + linear)."""
    N, D = z.shape
    m = (z.mean(0) + p.mean(0)) * 0.5
    zc, pc = z - m, p - m
    C = 2 * D * z_var
    Ce = C + 1e-7

    def dot(x, y):
        xh, xl = split(x); yh, yl = split(y)
        s = xh @ yh.t()
        if g1 >= 2: s = xl @ yh.t() + s
        if g1 == 3: s = xh @ yl.t() + s
        return s
    nz, npn = (zc * zc).sum(1), (pc * pc).sum(1)
    out64 = torch.zeros(N, D, dtype=torch.float64)
    O = torch.zeros(N, D)
    rs = torch.zeros(N)
    for (y, ny, sign) in ((zc, nz, 1.0), (pc, npn, -1.0)):
        dbar = (nz.mean() + ny.mean()).item()
        c0 = 1.0 / (Ce + dbar) ** 2
        c1 = -2.0 / (Ce + dbar) ** 3
        d2r = nz[:, None] + ny[None] - 2 * dot(zc, y)                 # rounded distances (fp32)
        c = 1.0 / (Ce + d2r) ** 2
        cpp = c - c0 - (c1 * (d2r - dbar) if order >= 1 else 0.0)      # what the tensor path carries
        cpp.fill_diagonal_(0.0)                                        # imq: no diagonal coefficient
        ch, cl = split(cpp); yh, yl = split(y)
        o = ch @ yh
        if g2 == 3: o = (cl @ yh + ch @ yl) + o
        O += sign * o
        rs += sign * cpp.sum(1)
        # analytic part in float64 with the EXACT distances
        zd, yd = zc.double(), y.double()
        d2 = (zd * zd).sum(1)[:, None] + (yd * yd).sum(1)[None] - 2 * zd @ yd.t()
        ca = torch.full_like(d2, c0) + (c1 * (d2 - dbar) if order >= 1 else 0.0)
        ca.fill_diagonal_(0.0)
        out64 += sign * (zd * ca.sum(1)[:, None] - ca @ yd)
    gsum = (zc * rs[:, None] - O).double() + out64
    return gsum * (-2 * C * 2)


if __name__ == "__main__":
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    for D, shift, sig in ((128, 0.25, 1.0), (128, 0.0, 1.0), (64, 0.0, 1.0), (16, 0.5, 1.0)):
        g0 = torch.Generator().manual_seed(1)
        z = sig * torch.randn(N, D, generator=g0) + shift; p = torch.randn(N, D, generator=g0)
        gref = ref64(z, p, 2.0)
        den = gref.abs().max()
        rows = []
        for order in (0, 1):
            for (g1, g2) in ((3, 3), (3, 1), (2, 1), (1, 1)):
                e = ((run(z, p, 2.0, g1, g2, order) - gref).abs().max() / den).item()
                rows.append("order %d g1=%d g2=%d: %.1e" % (order, g1, g2, e))
        print("N=%d D=%d shift=%.2f sig=%.1f imq | " % (N, D, shift, sig) + " | ".join(rows), flush=True)
