"""vaekit.io.DevicePrefetcher: batches come out in order and intact while the next copy is in flight."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))

pytestmark = pytest.mark.gpu


def test_prefetcher_order_and_contents():
    from vaekit.io import DevicePrefetcher
    dev = torch.device("cuda")
    batches = [{"z": torch.full((1 << 20,), float(i)).pin_memory(), "y": torch.arange(7, dtype=torch.int64).add(i).pin_memory()}
               for i in range(6)]
    pf = DevicePrefetcher(dev)
    pf.submit(batches[0])
    acc = torch.zeros((), device=dev, dtype=torch.float64)
    for i in range(6):
        d = pf.get()
        if i + 1 < 6:
            pf.submit(batches[i + 1])
        z = d["z"].requires_grad_(True)                     # what a step does to its input
        # a long kernel on the consumer stream: the slot must not be overwritten under it
        for _ in range(20):
            acc = acc + (z.detach().double() * 1.0).sum() / z.numel()
        assert d["z"].device.type == "cuda" and bool((d["z"] == float(i)).all())
        assert d["y"].tolist() == [i + k for k in range(7)]
    torch.cuda.synchronize()
    assert abs(acc.item() - 20 * sum(range(6))) < 1e-9
    with pytest.raises(RuntimeError, match="without a submitted batch"):
        pf.get()
    pf.submit(batches[0]); pf.submit(batches[1])
    with pytest.raises(RuntimeError, match="already in flight"):
        pf.submit(batches[2])
