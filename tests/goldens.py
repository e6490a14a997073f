"""Shared helpers to load tests/golden/*.npz and to evaluate one fixture with a
given implementation ("ops" object).  The same evaluation code is used for the
oracle (CPU, tests/test_oracle_golden.py) and for the CUDA product
(tests/test_gpu_golden.py) so the two are held to identical expectations."""
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name, device="cpu", dtype=torch.float32):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    out = {}
    for k in z.files:
        a = z[k]
        if k.startswith("hp_"):
            out[k] = a.item()
        elif a.dtype.kind == "f":
            out[k] = torch.from_numpy(a).to(device=device, dtype=dtype if k.startswith("in_") or k.startswith("noise_") else torch.float32)
        else:
            out[k] = torch.from_numpy(a).to(device)
    return out


def names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz"))


def leaf(t):
    return t.clone().detach().requires_grad_(True)


def rel_err(a, b):
    """max |a-b| / max(|b|) -- the 'relative' of the 1e-5 north-star tolerance for tensors."""
    a = torch.as_tensor(a).detach().double().cpu().reshape(-1)
    b = torch.as_tensor(b).detach().double().cpu().reshape(-1)
    den = b.abs().max().clamp_min(1e-30)
    return ((a - b).abs().max() / den).item()


def assert_close(a, b, tol, what=""):
    e = rel_err(a, b)
    assert e <= tol, f"{what}: rel err {e:.3e} > {tol:.1e}"
