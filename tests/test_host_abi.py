"""CPU-side checks of the boundary: libvaekit.so loads without a GPU, exports every symbol
include/vaekit.h declares (and nothing undeclared is bound), the host layer refuses CPU tensors
(no fallback), and the drop-in classes keep the reference's names/contracts."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
    import build as vk_build
    return vk_build.build()


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "vaekit.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vk_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(built):
    out = subprocess.run(["nm", "-D", "--defined-only", built], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (vk_[a-z0-9_]+)", out))
    decl = header_symbols()
    assert decl, "no declarations parsed"
    missing = [s for s in decl if s not in exported]
    assert not missing, missing


def test_ctypes_binding_covers_header(built):
    import vaekit
    assert sorted(vaekit.declared_symbols()) == header_symbols()
    lib = vaekit.lib()
    assert lib.vk_version() >= 100
    assert lib.vk_workspace_bytes() > 256


def test_no_cpu_fallback(built):
    from vaekit import functional as F
    with pytest.raises(RuntimeError, match="CUDA"):
        F.kl_rows(torch.zeros(2, 4), torch.zeros(2, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        F.iw_bound(torch.zeros(2, 3, 8), torch.zeros(2, 8), torch.zeros(2, 3), 0.1)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "pytorch-vae_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                bad = re.findall(r"^\s*(?:from|import)\s+\S*oracle|#include.*oracle|sys\.path.*oracle", txt, flags=re.M)
                assert not bad, (os.path.join(dp, f), bad)


def test_registry_and_constructor_contracts(built):
    from vaekit.models import vae_models
    m = vae_models["BetaVAE"](3, 10, loss_type="H", beta=10, some_unused_kwarg=1)
    assert m.num_iter == 0 and m.C_max.shape == (1,)
    m = vae_models["MIWAE"](3, 10)
    assert (m.num_samples, m.num_estimates) == (5, 5)
    assert vae_models["LogCoshVAE"](3, 10).alpha == 100.0
    hd = [16, 32]
    vae_models["VanillaVAE"](3, 4, hidden_dims=hd)
    assert hd == [32, 16]          # the reference reverses the caller's list in place (vanilla_vae.py:45)
    m = vae_models["InfoVAE"](3, 16, alpha=-9.0, beta=10.5, reg_weight=110)
    assert (m.alpha, m.beta, m.reg_weight, m.kernel_type, m.z_var) == (-9.0, 10.5, 110, "imq", 2.0)
    with pytest.raises(AssertionError, match="alpha must be negative or zero"):
        vae_models["InfoVAE"](3, 16, alpha=0.5)          # info_vae.py:27


def test_step_report_packs_one_copy(built):
    """SURVEY 8(f) rank 3 host logic: loss dict (+ per-image values) -> Python floats through one concatenated tensor."""
    from vaekit import functional as F
    d = {"loss": torch.tensor([0.5]), "Reconstruction_Loss": torch.tensor(0.25), "KLD": torch.tensor(-1.0)}
    vals, per = F.step_report(d, torch.tensor([1.0, 2.0, 3.0]))
    assert vals == {"loss": 0.5, "Reconstruction_Loss": 0.25, "KLD": -1.0} and per == [1.0, 2.0, 3.0]
    with pytest.raises(RuntimeError, match="CUDA"):
        F.per_image_mse(torch.zeros(2, 3, 4, 4), torch.zeros(2, 3, 4, 4))


def test_last_model_host_parts(built):
    """Torch-side pieces of the MSSIMVAE / DFCVAE / VampVAE drop-ins that need no GPU: the MS-SSIM module against the
    reference fixture (quirks of mssim_vae.py:80-82, 206-209, 282 included), the DFCVAE extractor hook and constructor
    contracts.  Their loss_function kernels are covered by tests/test_gpu_next_models.py."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from goldens import assert_close, leaf, load
    from vaekit.models import vae_models
    from vaekit.models.mssim_vae import MSSIM
    g = load("mssimvae")
    m = vae_models["MSSIMVAE"](3, 16)
    assert m.mssim_loss.normalize is True and m.mssim_loss.size_average is True and m.mssim_loss.window_size == 11
    assert vae_models["MSSIMVAE"](3, 16, size_average=False).mssim_loss.normalize is False      # lands in `normalize`
    r = leaf(g["in_recons"])
    out = m.mssim_loss(r, g["in_input"])
    assert_close(out, g["out_Reconstruction_Loss"], 1e-6, "1 - MS-SSIM")
    out.backward()
    w = MSSIM().gaussian_window(11, 1.5)
    assert abs(float(w.sum()) - 1.0) < 1e-6 and w[0] > w[5]            # exp(+d^2): heaviest at the edges

    feats = torch.nn.Sequential(*[torch.nn.Conv2d(3, 3, 1) if i % 3 == 0 else torch.nn.ReLU() for i in range(44)])
    net = torch.nn.Module()
    net.features = feats
    d = vae_models["DFCVAE"](3, 16, feature_network=net)
    assert (d.alpha, d.beta) == (1, 0.5) and not d.feature_network.training
    assert all(not p.requires_grad for p in d.feature_network.parameters())
    f = d.extract_features(torch.rand(2, 3, 8, 8))
    assert len(f) == 4 and len(d.extract_features(torch.rand(2, 3, 8, 8), ['0', '1'])) == 2

    v = vae_models["VampVAE"](3, 16, num_components=12)
    assert v.pseudo_input.shape == (12, 12) and not v.pseudo_input.requires_grad
    assert "pseudo_input" not in v.state_dict() and v.embed_pseudo[0].out_features == 12288
    pm, pv = v.prior_components(torch.rand(2, 3, 64, 64))
    assert pm.shape == (12, 16) and pv.shape == (12, 16)


def test_bench_host_helpers_are_best_effort():
    """bench.py's NUMA binding must never take a run down (no NVML / no GPU here: it returns None and leaves the affinity
    alone), and the prefetcher refuses misuse before touching a device."""
    import os
    sys.path.insert(0, ROOT)
    import bench
    before = os.sched_getaffinity(0)
    got = bench.bind_to_gpu_numa_node(0)
    assert got is None or (isinstance(got, int) and got > 0)
    if got is None:
        assert os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)
    src = open(os.path.join(ROOT, "pytorch-vae_b200", "vaekit", "io.py")).read()
    assert "oracle" not in src and "wait_event" in src                      # plumbing only: streams and events
