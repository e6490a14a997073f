#!/usr/bin/env python
"""Benchmark of the latent + loss hot path (BASELINE.json metric: samples/s fwd+bwd; % of HBM / tensor roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload all|<name>] [--impl reference]

ONE invocation times every BASELINE.json config (and BetaTC) and prints ONE JSON line on stdout (rank 0):

  top level   = BASELINE configs[2], the config the north star puts its multi-GPU target on: WAE-MMD RBF, z batch
                65536 x 128, value + d/dz through the drop-in model API.  Under torchrun it is the row-sharded form
                (all-gather of z, 128-aligned row blocks per rank): "scaling": "strong".
  "workloads" = {name: {value, ms_per_step, roofline, cpu_baseline, e2e, clocks, ...}} for the other configs:
                wae_mmd_imq, iwae, miwae (configs[1]), vanilla_b64 (configs[0] as BASELINE words it: full model(x) +
                loss_function + backward at batch 64), vanilla (the same loss at streaming size), vq (configs[3]),
                swae (configs[4]), betatc (SURVEY 8(d) extra).  Under torchrun: the strong-scaled sharded MMD / SWD /
                BetaTC paths plus the batch-local ops as weak-scaled replicas.

Every workload: W (>= 3) warm-up steps, K timed steps between two CUDA events on the launching stream (max over ranks),
then the same K steps again with CUDA events around every library call for the roofline of the dominant call, then the
end-to-end leg (pinned host inputs copied in and the loss read back every step), then -- rank 0, one GPU only -- the
reference's CPU formulation (oracle port) on a bounded sample.  `--impl reference` times that CPU path for the top-level
workload with the same K / W.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "pytorch-vae_b200"),):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

P_IMG = 3 * 64 * 64
METRIC = "latent+loss samples/s fwd+bwd"
HEADLINE = "wae_mmd_rbf"


def _peaks_file():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return None


def hbm_peak():
    pk = _peaks_file()
    if pk and "hbm_gbs" in pk:
        return float(pk["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def bf16_peak():
    pk = _peaks_file()
    if pk and "bf16_tflops_sustained" in pk:
        return float(pk["bf16_tflops_sustained"]), "MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)"
    return 1400.0, "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML every few ms while the timed region runs."""

    def __init__(self, index, period_s=0.004):
        self.index, self.period, self.samples, self.reasons = index, period_s, [], set()
        self.max_mhz, self._stop, self.thread, self.err = None, threading.Event(), None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()
        except Exception as e:          # no NVML: report it rather than invent numbers
            self.err = repr(e)
        return self

    def _run(self):
        nv = self.nv
        bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, b in bits.items():
                    if r & b:
                        self.reasons.add(k)
            except Exception as e:
                self.err = repr(e)
                return
            time.sleep(self.period)

    def stop(self):
        self._stop.set()
        if self.thread is not None:
            self.thread.join(timeout=2)
        sm = sorted(self.samples)
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
               "reasons": sorted(self.reasons), "samples": len(sm)}
        if self.err:
            out["error"] = self.err
        return out


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def oracle():
    """The CPU oracle: test infrastructure, used here only for the cpu_baseline / --impl reference legs."""
    p = os.path.join(ROOT, "oracle")
    if p not in sys.path:
        sys.path.insert(0, p)
    import vae_oracle
    return vae_oracle


# =====================================================================================================================
# workloads
# =====================================================================================================================
class Workload:
    """One hot-path op at a BASELINE.json size.  make() builds synthetic inputs, step() runs forward + backward through
    the public (drop-in model) API, cpu_make() / cpu_step() are the oracle port of the reference's formulation on a
    bounded sample."""
    scaling = "weak"            # replicas: every rank owns its own batch, no data-path collective
    bound = "hbm"
    dominant = None             # library call whose roofline is reported; None = the call with the largest share
    parallelism = "batch-local replicas, no data-path collective"
    graphed = False
    l2_policy = "inputs larger than L2, no flush"
    ncu_traffic = {}

    def units(self, world):
        return self.N * (world if self.scaling == "weak" else 1)

    def model(self, dev):
        raise NotImplementedError

    def extras(self, model, t, args, dev):
        return {}


# ---- configs[1]: IWAE / MIWAE, configs[0] at streaming size -------------------------------------------------------
class StreamWorkload(Workload):
    """reparameterize + loss_function + backward on given tensors (K1 + K2 / K3)."""

    def __init__(self, name):
        cfg = {"iwae": dict(B=4096, G=1, K=5, M_N=0.00025,
                            desc="IWAE K=5, batch 4096, latent 128, 3x64x64: reparameterize + importance-weighted bound (BASELINE configs[1])"),
               "miwae": dict(B=4096, G=3, K=5, M_N=0.00025,
                             desc="MIWAE K=5 M=3, batch 4096, latent 128, 3x64x64 (BASELINE configs[1])"),
               "vanilla": dict(B=65536, G=0, K=0, M_N=0.005,
                               desc="VanillaVAE reparameterize + loss_function on given tensors, batch 65536, latent 128, 3x64x64 "
                                    "(BASELINE configs[0] at streaming size)")}[name]
        self.name, self.N, self.G, self.K, self.M_N, self.desc, self.D = name, cfg["B"], cfg["G"], cfg["K"], cfg["M_N"], cfg["desc"], 128
        self.cpu_sample = 1024 if self.K else 4096
        self.ncu_traffic = {"iw_bound_bwd": (1.208158e9 + 952.969216e6, "profiles/ncu_iw_bound_r01a_summary.csv"),
                            "iw_bound_fwd": (1.208144e9 + 4.415744e6, "profiles/ncu_iw_bound_r01a_summary.csv")} if name == "iwae" else {}

    def model(self, dev):
        from vaekit.models import vae_models
        if self.name == "iwae":
            return vae_models["IWAE"](3, self.D, num_samples=5)
        if self.name == "miwae":
            return vae_models["MIWAE"](3, self.D, num_samples=5, num_estimates=3)
        return vae_models["VanillaVAE"](3, self.D)

    def _ss(self, B):
        if self.K == 0:
            return (B,)
        return (B, self.G, self.K) if self.G > 1 else (B, self.K)

    def make(self, dev, rank, world, n=None):
        B = n or self.N
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        ss = self._ss(B)
        return {"recons": torch.tanh(torch.randn(*ss, 3, 64, 64, device=dev, generator=g)),     # decoder ends in Tanh
                "input": torch.rand(B, 3, 64, 64, device=dev, generator=g),                       # ToTensor range
                "mu0": torch.randn(B, self.D, device=dev, generator=g),
                "lv0": 0.5 * torch.randn(B, self.D, device=dev, generator=g) - 1.0,
                "gz": 1e-3 * torch.randn(*ss, self.D, device=dev, generator=g)}                   # upstream grad of z

    def _expand(self, mu0, lv0):
        if self.K == 0:
            return mu0, lv0
        if self.G > 1:
            shp = (-1, self.G, self.K, -1)
            return mu0.view(mu0.size(0), 1, 1, -1).expand(*shp), lv0.view(lv0.size(0), 1, 1, -1).expand(*shp)
        return mu0.unsqueeze(1).expand(-1, self.K, -1), lv0.unsqueeze(1).expand(-1, self.K, -1)

    def step(self, m, t, world=1):
        recons = t["recons"].requires_grad_(True)
        mu0 = t["mu0"].requires_grad_(True)
        lv0 = t["lv0"].requires_grad_(True)
        recons.grad = mu0.grad = lv0.grad = None
        mu, lv = self._expand(mu0, lv0)
        z = m.reparameterize(mu, lv)
        if self.K == 0:
            d = m.loss_function(recons, t["input"], mu, lv, M_N=self.M_N)
        else:
            d = m.loss_function(recons, t["input"], mu, lv, z, None, M_N=self.M_N)
        torch.autograd.backward([d["loss"], z], [None, t["gz"]])
        return {k: v.detach() for k, v in d.items()}

    def work(self, world):
        B, D, GK = self.N, self.D, max(self.G, 1) * self.K
        return {"iw_bound_fwd": B * (GK * P_IMG + P_IMG) * 4 + B * GK * 8,
                "iw_bound_bwd": B * (2 * GK * P_IMG + P_IMG) * 4 + B * GK * 12,
                "elbo_fwd": B * 2 * P_IMG * 4 + B * 4, "elbo_bwd": B * 3 * P_IMG * 4 + B * 4,
                "reparam_kl_fwd": B * 2 * D * 4 + B * max(GK, 1) * (D * 4 + 4),
                "reparam_kl_bwd": B * 2 * D * 4 + B * max(GK, 1) * (3 * D * 4 + 4)}, "GB/s"

    def cpu_make(self):
        return {k: v for k, v in self.make(torch.device("cpu"), 0, 1, n=self.cpu_sample).items()}, self.cpu_sample, \
            "batch %d of the %d-sample workload (cost per sample does not depend on the batch)" % (self.cpu_sample, self.N)

    def cpu_step(self, t):
        """iwae.py:121-160 / miwae.py:124-164 / vanilla_vae.py:119-146 as the reference composes them."""
        O = oracle()
        recons = t["recons"].detach().requires_grad_(True)
        mu0 = t["mu0"].detach().requires_grad_(True)
        lv0 = t["lv0"].detach().requires_grad_(True)
        if self.K == 0:
            mu, lv = mu0, lv0
        elif self.G > 1:
            mu = mu0.repeat(self.G, self.K, 1, 1).permute(2, 0, 1, 3)
            lv = lv0.repeat(self.G, self.K, 1, 1).permute(2, 0, 1, 3)
        else:
            mu = mu0.repeat(self.K, 1, 1).permute(1, 0, 2)
            lv = lv0.repeat(self.K, 1, 1).permute(1, 0, 2)
        z = O.reparameterize(mu, lv, torch.randn_like(mu))
        if self.K == 0:
            d = O.vanilla_loss(recons, t["input"], mu, lv, self.M_N)
        else:
            d = O.iw_bound_loss(recons, t["input"], mu, lv, self.M_N)
        torch.autograd.backward([d["loss"], z], [None, t["gz"]])


# ---- configs[0] as BASELINE.json words it ----------------------------------------------------------------------------
class VanillaB64(Workload):
    """tests/test_vae.py:11-28 + configs/vae/vae.yaml: VanillaVAE(3, 128), batch 64 of 64x64 RGB, the full model(x) +
    loss_function + backward.  The hot path moves 16 MB per step (2.5 us of HBM time): the step is launch-latency-bound
    (SURVEY 0.10), so it is reported as us/step and launches/step, eager and replayed from ONE CUDA graph (K1 reads its
    Philox key from device memory, so the replay draws fresh noise)."""
    graphed = True
    l2_policy = "working set fits L2 by construction (16 MB): latency-bound config, reported as us/step"
    dominant = "elbo_bwd"

    def __init__(self):
        self.name, self.N, self.D = "vanilla_b64", 64, 128
        self.desc = ("VanillaVAE(3, 128) full model(x) + loss_function + backward, batch 64, 64x64 RGB (BASELINE configs[0]; "
                     "encoder / decoder convs are torch modules, out of scope but inside the step)")

    def model(self, dev):
        from vaekit.models import vae_models
        torch.manual_seed(1265)
        return vae_models["VanillaVAE"](3, self.D).to(dev)

    def make(self, dev, rank, world, n=None):
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        return {"x": torch.rand(n or self.N, 3, 64, 64, device=dev, generator=g)}

    def step(self, m, t, world=1):
        for p in m.parameters():
            p.grad = None
        out = m(t["x"])
        d = m.loss_function(*out, M_N=0.005)
        d["loss"].backward()
        return {"loss": d["loss"].detach()}

    def work(self, world):
        B, D = self.N, self.D
        return {"elbo_fwd": B * 2 * P_IMG * 4, "elbo_bwd": B * 3 * P_IMG * 4, "reparam_kl_fwd": B * 3 * D * 4,
                "reparam_kl_bwd": B * 5 * D * 4}, "GB/s"

    def extras(self, m, t, args, dev):
        """The hot path alone (reparameterize + loss_function + backward on given tensors, no convs), eager and graphed."""
        from vaekit.graphs import GraphedStep
        from vaekit import _lib as L
        sw = StreamWorkload("vanilla")
        sw.N = self.N
        tt = sw.make(dev, 0, 1, n=self.N)
        mm = sw.model(dev)

        def timed(fn, steps):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n0 = L.LAUNCHES
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps, (L.LAUNCHES - n0) / steps

        for _ in range(3):
            sw.step(mm, tt)
        eager_ms, launches = timed(lambda: sw.step(mm, tt), max(args.steps, 20))
        gs = GraphedStep(lambda inp: sw.step(mm, inp)["loss"], tt, rng=True)
        graph_ms, _ = timed(gs, max(args.steps, 20))
        alg = B_bytes = self.N * (5 * P_IMG * 4 + 8 * self.D * 4)
        peak, _ = hbm_peak()
        return {"hot_path_only": {"what": "reparameterize + loss_function + backward on given [64, ...] tensors (no convs)",
                                  "eager_us_per_step": round(1e3 * eager_ms, 2), "graph_us_per_step": round(1e3 * graph_ms, 2),
                                  "vaekit_launches_per_step": launches, "algorithmic_bytes_per_step": B_bytes,
                                  "graph_GBps": round(alg / graph_ms / 1e6, 1), "graph_frac_of_hbm": round(alg / graph_ms / 1e6 / peak, 4),
                                  "hbm_time_us_at_peak": round(alg / peak / 1e3, 2)}}

    def cpu_make(self):
        torch.manual_seed(1265)
        from vaekit.models import vae_models
        m = vae_models["VanillaVAE"](3, self.D)                       # conv stacks are plain torch modules: they run on the CPU
        return {"x": torch.rand(self.N, 3, 64, 64), "m": m}, self.N, "the full config: batch 64, model(x) + loss_function + backward"

    def cpu_step(self, t):
        """vanilla_vae.py:119-146 with the reference's torch ops: conv stacks + oracle reparameterize / loss."""
        O = oracle()
        m = t["m"]
        for p in m.parameters():
            p.grad = None
        mu, lv = m.encode(t["x"])
        z = O.reparameterize(mu, lv, torch.randn_like(mu))
        O.vanilla_loss(m.decode(z), t["x"], mu, lv, 0.005)["loss"].backward()


# ---- configs[2]: WAE-MMD ---------------------------------------------------------------------------------------------
class WaeMmd(Workload):
    scaling = "strong"          # fixed global batch, rows sharded after an all-gather of z
    bound = "tensor"
    dominant = "mmd_fwd"
    parallelism = "row-sharded after an NCCL all-gather of z (33.5 MB), 128-aligned row block per rank, all-reduce of 3 float64 sums"
    l2_policy = "tensor-bound: operand images (67 MB) are L2-resident by design; nothing to flush"

    def __init__(self, kt):
        self.kt, self.N, self.D = kt, 65536, 128
        self.name = "wae_mmd_" + kt
        self.reg = 5000.0 if kt == "rbf" else 100.0
        self.desc = "WAE-MMD %s kernel, z batch 65536x128, value + d/dz in one tcgen05 Gram sweep (BASELINE configs[2])" % kt.upper()
        self.cpu_rows = 8
        if kt == "rbf":      # dram__bytes_read + write of mmd_main_kernel<2,1,0>, the launch that is 91 % of the call
            self.ncu_traffic = {"mmd_fwd": (74.697472e6 + 39.711232e6, "profiles/ncu_mmd_rbf_r02z_summary.txt")}

    def model(self, dev):
        from vaekit.models import vae_models
        return vae_models["WAE_MMD"](3, self.D, reg_weight=self.reg, kernel_type=self.kt)

    def make(self, dev, rank, world, n=None):
        n = n or self.N // world
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        return {"z": torch.randn(n, self.D, device=dev, generator=g) + 0.25}

    def step(self, m, t, world=1):
        z = t["z"].requires_grad_(True)
        z.grad = None
        n_glob = z.shape[0] * world
        reg = m.reg_weight / (n_glob * (n_glob - 1))
        if world > 1:
            from vaekit import dist as vd
            mmd, _ = vd.sharded_mmd(z, reg, m.kernel_type, m.z_var)
        else:
            mmd = m.compute_mmd(z, reg)
        mmd.backward()
        return {"MMD": mmd.detach()}

    def work(self, world):
        return {"mmd_fwd": 10.0 * self.N * self.N * self.D / world}, "TFLOP/s"

    def cpu_make(self):
        t = self.make(torch.device("cpu"), 0, 1, n=self.N)
        t["prior"] = torch.randn(self.N, self.D)
        return t, self.cpu_rows, ("a block of %d rows x all 65536 columns of the three kernel matrices in the reference's own "
                                  "broadcast formulation at the FULL N (cost = rows/N of the step; the full [N,N,D] temporaries "
                                  "need 2.2 TB)" % self.cpu_rows)

    def cpu_step(self, t):
        O = oracle()
        z = t["z"].detach().requires_grad_(True)
        O.mmd_row_block(z, t["prior"], self.cpu_rows, self.kt, 2.0).backward()


# ---- configs[3]: VQ --------------------------------------------------------------------------------------------------
class VqWorkload(Workload):
    dominant = "vq_fwd"
    graphed = True

    def __init__(self):
        self.name, self.N, self.D, self.K, self.H = "vq", 1024, 64, 512, 16
        self.desc = "VQ-VAE VectorQuantizer, codebook 512x64, 16x16 latent grid, batch 1024: argmin + straight-through + commitment loss (BASELINE configs[3])"
        self.cpu_sample = 128

    def model(self, dev):
        from vaekit.models.vq_vae import VectorQuantizer
        torch.manual_seed(1265)
        m = VectorQuantizer(self.K, self.D, 0.25)
        with torch.no_grad():
            m.embedding.weight.normal_()
        return m.to(dev)

    def make(self, dev, rank, world, n=None):
        n = n or self.N
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        return {"latents": torch.nn.functional.leaky_relu(torch.randn(n, self.D, self.H, self.H, device=dev, generator=g)),
                "gq": 1e-3 * torch.randn(n, self.D, self.H, self.H, device=dev, generator=g)}

    def step(self, m, t, world=1):
        x = t["latents"].requires_grad_(True)
        x.grad = None
        m.embedding.weight.grad = None
        q, vq_loss = m(x)
        torch.autograd.backward([q, vq_loss], [t["gq"], None])
        return {"VQ_Loss": vq_loss.detach()}

    def work(self, world):
        R = self.N * self.H * self.H
        fwd = 2 * R * self.D * 4 + R * 4 + self.K * self.D * 4
        bwd = 4 * R * self.D * 4 + R * 4 + 2 * self.K * self.D * 4
        return {"vq_fwd": float(fwd), "vq_bwd": float(bwd)}, "GB/s"

    def extras(self, m, t, args, dev):
        """Index mismatch rate over all 262 144 rows by gap bucket (north star: bit-exact wherever the reference's fp32
        distance gap exceeds 1e-5 relative; the rate is reported).  The checker is torch on the GPU: the reference's
        own expression (vq_vae.py:32-37) in fp32 and in float64."""
        from vaekit import functional as F
        x, e = t["latents"].detach(), m.embedding.weight.detach()
        _, _, idx = F.vector_quantize(x, e, 0.25)
        flat = x.permute(0, 2, 3, 1).reshape(-1, self.D)
        tf32 = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        try:
            d32 = torch.sum(flat ** 2, dim=1, keepdim=True) + torch.sum(e ** 2, dim=1) - 2 * torch.matmul(flat, e.t())
        finally:
            torch.backends.cuda.matmul.allow_tf32 = tf32
        i32 = torch.argmin(d32, dim=1)
        best2 = torch.topk(d32, 2, dim=1, largest=False).values
        gap = (best2[:, 1] - best2[:, 0]) / best2[:, 1].abs().clamp_min(1e-30)
        del d32
        f64, e64 = flat.double(), e.double()
        i64 = torch.argmin(torch.sum(f64 ** 2, dim=1, keepdim=True) + torch.sum(e64 ** 2, dim=1) - 2 * f64 @ e64.t(), dim=1)
        ours = idx.long()
        buckets = {}
        for name, lo, hi in (("gap>1e-5", 1e-5, float("inf")), ("1e-6<gap<=1e-5", 1e-6, 1e-5), ("0<gap<=1e-6", 0.0, 1e-6),
                             ("exact_tie", -1.0, 0.0)):
            sel = (gap > lo) & (gap <= hi)
            n = int(sel.sum())
            buckets[name] = {"rows": n, "mismatch_vs_reference_fp32": int((ours[sel] != i32[sel]).sum()),
                             "mismatch_vs_float64": int((ours[sel] != i64[sel]).sum()),
                             "reference_fp32_vs_float64": int((i32[sel] != i64[sel]).sum())}
        return {"index_mismatch": {"rows": int(ours.numel()), "overall_vs_reference_fp32": int((ours != i32).sum()),
                                   "overall_vs_float64": int((ours != i64).sum()),
                                   "reference_fp32_vs_float64": int((i32 != i64).sum()), "by_relative_gap": buckets,
                                   "checker": "torch on the GPU: vq_vae.py:32-37 in fp32 (TF32 off) and in float64"}}

    def cpu_make(self):
        t = self.make(torch.device("cpu"), 0, 1, n=self.cpu_sample)
        t["cb"] = torch.randn(self.K, self.D)
        return t, self.cpu_sample, "batch %d of 1024 images (cost per image does not depend on the batch)" % self.cpu_sample

    def cpu_step(self, t):
        O = oracle()
        x = t["latents"].detach().requires_grad_(True)
        cb = t["cb"].detach().requires_grad_(True)
        q, l, _ = O.vq_forward(x, cb, 0.25)
        torch.autograd.backward([q, l], [t["gq"], None])


# ---- configs[4]: SWAE ------------------------------------------------------------------------------------------------
class SwaeWorkload(Workload):
    scaling = "strong"          # fixed global batch, every stage sharded
    dominant = "swd_fwd"
    dominant_sharded = "swd_segments_fwd"
    l2_policy = "sort buffers (1.7 GB per pass) larger than L2, no flush"

    def __init__(self):
        self.name, self.N, self.D, self.S = "swae", 1 << 20, 128, 200
        self.desc = "SWAE sliced-Wasserstein, 200 projections, batch 1048576x128: projection GEMM + per-projection radix sort (BASELINE configs[4])"
        self.cpu_proj = 20
        # dram__bytes_read.sum + dram__bytes_write.sum over the 19 kernels of one vk_swd_fwd call at this size (r01v)
        self.ncu_traffic = {"swd_fwd": (30.822819e9, "profiles/ncu_swae_r01v_summary.txt")}
        self.parallelism = ("rows and projections both sharded: every rank projects its rows on all 200 projections, sorts its "
                            "200/world projections; both exchanges fused into the tcgen05 GEMMs over peer memory (%s)"
                            % os.environ.get("VAEKIT_SWD_EXCHANGE", "auto"))

    def model(self, dev):
        from vaekit.models import vae_models
        return vae_models["SWAE"](3, self.D, reg_weight=100, num_projections=self.S)

    def make(self, dev, rank, world, n=None):
        n = n or self.N // world
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        return {"z": 1.5 * torch.randn(n, self.D, device=dev, generator=g) + 0.1}

    def step(self, m, t, world=1):
        z = t["z"].requires_grad_(True)
        z.grad = None
        n_glob = z.shape[0] * world
        reg = m.reg_weight / (n_glob * (n_glob - 1))
        if world > 1:
            from vaekit import dist as vd
            proj = vd.shared_projections(m, self.D, self.S, z.device)
            swd = vd.sharded_swd(z, proj, m.p, reg, exchange=os.environ.get("VAEKIT_SWD_EXCHANGE", "auto"))
        else:
            swd = m.compute_swd(z, m.p, reg)
        swd.backward()
        return {"SWD": swd.detach()}

    def work(self, world):
        # radix-sort traffic model of SURVEY 8(d): 27 424 B/sample; sharded runs time the sort + loss stage on its own (120 of
        # the model's 132 bytes per (sample, projection): key+index sort 68, key-only sort 36, loss read 8, scatter 8)
        return {"swd_fwd": 27424.0 * self.N / world, "swd_segments_fwd": 120.0 * self.N * self.S / world}, "GB/s"

    def extras(self, m, t, args, dev):
        return {"compulsory_floor": {"bytes_per_sample": 1024, "what": "z read + g_z write only (SURVEY 8(d): everything else on chip)"}}

    def cpu_make(self):
        O = oracle()
        t = self.make(torch.device("cpu"), 0, 1, n=self.N)
        t["prior"] = torch.randn(self.N, self.D)
        torch.manual_seed(1265)
        t["proj"] = O.random_projections(self.D, self.S)[:self.cpu_proj].contiguous()
        return t, self.N * self.cpu_proj / self.S, ("%d of the 200 projections at the FULL batch 1048576 (cost is linear in the "
                                                    "projection count; the prior draw is outside the timed region)" % self.cpu_proj)

    def cpu_step(self, t):
        O = oracle()
        z = t["z"].detach().requires_grad_(True)
        n = z.shape[0]
        O.swd(z, t["prior"], t["proj"], 2.0, 100.0 / (n * (n - 1))).backward()


# ---- SURVEY 8(d) extra: BetaTC -----------------------------------------------------------------------------------------
class BetaTcWorkload(Workload):
    scaling = "strong"          # fixed global batch, anchor rows sharded after an all-gather of z, mu, log_var
    bound = "mufu"
    dominant, dominant_sharded = "tc_bwd", "tc_bwd_rows"
    parallelism = "anchor rows sharded after NCCL all-gathers of z, mu, log_var; all-reduce of the 4 batch means"
    l2_policy = "MUFU-bound: operands (6 MB) are L2-resident by design"

    def __init__(self):
        self.name, self.N, self.D = "betatc", 4096, 128
        self.desc = "BetaTC-VAE latent terms (log_q_z, per-dimension LSE, MI / TC / KLD), batch 4096, latent 128 (SURVEY 8(d) extra)"
        self.cpu_rows = 64

    def model(self, dev):
        from vaekit.models import vae_models
        return vae_models["BetaTCVAE"](3, self.D).eval()

    def make(self, dev, rank, world, n=None):
        n = n or self.N // world
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        mu = torch.randn(n, self.D, device=dev, generator=g)
        lv = 0.3 * torch.randn(n, self.D, device=dev, generator=g)
        return {"mu": mu, "lv": lv}

    def step(self, m, t, world=1):
        from vaekit import functional as F
        mu, lv = t["mu"].requires_grad_(True), t["lv"].requires_grad_(True)
        mu.grad = lv.grad = None
        z = m.reparameterize(mu, lv)
        if world > 1:
            from vaekit import dist as vd
            latent, terms = vd.sharded_betatc_terms(z, mu, lv, 0.00025, m.alpha, m.beta, m.gamma)
        else:
            latent, terms = F.betatc_terms(z, mu, lv, 0.00025, m.alpha, m.beta, m.gamma)
        latent.backward()
        return {"latent": latent.detach()}

    def work(self, world):
        e = float(self.N) * self.N * self.D
        return {"tc_fwd": e, "tc_bwd": 2.0 * e, "tc_fwd_rows": e / world, "tc_bwd_rows": 2.0 * e / world}, "Gexp/s"

    def cpu_make(self):
        t = self.make(torch.device("cpu"), 0, 1, n=self.N)
        t["eps"] = torch.randn(self.N, self.D)
        return t, self.cpu_rows, ("%d anchor rows x all 4096 partners of the reference's [B,B,D] log-density tensor at the FULL "
                                  "batch (cost = rows/B of the step; the full tensor is 8.6 GB per temporary)" % self.cpu_rows)

    def cpu_step(self, t):
        O = oracle()
        mu, lv = t["mu"].detach().requires_grad_(True), t["lv"].detach().requires_grad_(True)
        z = O.reparameterize(mu, lv, t["eps"])
        O.betatc_row_block(mu, lv, z, self.cpu_rows, 0.00025).backward()


WORKLOADS = {"wae_mmd_rbf": lambda: WaeMmd("rbf"), "wae_mmd_imq": lambda: WaeMmd("imq"),
             "iwae": lambda: StreamWorkload("iwae"), "miwae": lambda: StreamWorkload("miwae"),
             "vanilla_b64": VanillaB64, "vanilla": lambda: StreamWorkload("vanilla"),
             "vq": VqWorkload, "swae": SwaeWorkload, "betatc": BetaTcWorkload}
ORDER_1GPU = ["wae_mmd_rbf", "wae_mmd_imq", "iwae", "miwae", "vanilla_b64", "vanilla", "vq", "swae", "betatc"]
ORDER_NGPU = ["wae_mmd_rbf", "wae_mmd_imq", "swae", "betatc", "iwae", "vq"]


# =====================================================================================================================
# CPU arm: the reference's formulation (oracle port) on the host cores
# =====================================================================================================================
def cpu_time(wl, iters, warmup):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    t, units, sample = wl.cpu_make()
    for _ in range(warmup):
        wl.cpu_step(t)
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        wl.cpu_step(t)
        ts.append(time.perf_counter() - t0)
    return units * len(ts) / sum(ts), cores, ts, sample, units


def cpu_baseline(wl, budget_s=8.0):
    """Bounded sample: one warm-up, then as many iterations as fit the budget (at least 2, at most 5)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    t, units, sample = wl.cpu_make()
    t0 = time.perf_counter()
    wl.cpu_step(t)
    first = time.perf_counter() - t0
    iters = max(2, min(5, int(budget_s / max(first, 1e-3))))
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        wl.cpu_step(t)
        ts.append(time.perf_counter() - t0)
    v = units * len(ts) / sum(ts)
    return {"value": round(v, 3), "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%s; %d iterations after 1 warm-up, torch CPU %d threads; oracle/vae_oracle.py = the reference's torch ops"
                      % (sample, iters, cores), "s_per_iteration": round(sum(ts) / len(ts), 3)}


def run_reference_arm(args):
    """--impl reference: the reference's CPU formulation of the top-level workload (oracle port of its torch code -- the
    Python reference itself cannot travel to the GPU box and has no native code to compile) on all host cores, each step
    a bounded sample of that workload."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    name = HEADLINE if args.workload == "all" else args.workload
    wl = WORKLOADS[name]()
    v, cores, ts, sample, units = cpu_time(wl, args.steps, max(1, min(args.warmup, 2)))
    line = {"impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * sum(ts) / len(ts), 3),
            "higher_is_better": True, "scaling": wl.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.desc, "sample": sample, "samples_per_step": units},
            "cpu_baseline": {"value": round(v, 3), "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(v, 3), "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# =====================================================================================================================
# GPU arm
# =====================================================================================================================
def bind_to_gpu_numa_node(local_rank):
    """Pins this rank's process to the CPUs NVML reports as local to its GPU BEFORE any pinned host memory is allocated
    (first touch puts the pages on that node): on an 8-GPU box the ranks otherwise pull their batches through one
    socket's memory and the inter-socket link.  Best effort: returns the CPU count it bound to, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local_rank]) if vis and vis.split(",")[local_rank].isdigit() else local_rank
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


class Ctx:
    def __init__(self, args):
        import torch.distributed as dist
        self.dist = dist
        self.rank, self.world, self.local = dist_env()
        self.numa = bind_to_gpu_numa_node(self.local) if self.world > 1 else None
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.args = args

    def barrier(self):
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.world > 1:
            t = torch.tensor([ms], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = t.item()
        return ms


def timed(ctx, fn, steps, log_events=False):
    """K calls of fn between two CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks."""
    from vaekit import _lib as L
    ctx.barrier()
    if log_events:
        L.EVENT_LOG = []
    n0 = L.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        r = fn()
    e1.record()
    ctx.barrier()
    log, L.EVENT_LOG = L.EVENT_LOG, None
    return ctx.max_over_ranks(e0.elapsed_time(e1)), log, L.LAUNCHES - n0, r


def roofline(wl, world, log):
    work, unit = wl.work(world)
    acc = {}
    for name, a, b in log:
        acc.setdefault(name, []).append(a.elapsed_time(b))
    calls = {k: {"ms": round(sum(v) / len(v), 4), "calls": len(v)} for k, v in acc.items()}
    tot = sum(k["ms"] * k["calls"] for k in calls.values()) or 1.0
    for k in calls:
        calls[k]["share_of_library_time"] = round(calls[k]["ms"] * calls[k]["calls"] / tot, 4)
        if k in work:
            calls[k]["algorithmic_work"] = work[k]
    dom = wl.dominant if wl.dominant in calls else getattr(wl, "dominant_sharded", None)
    if dom not in calls:
        cand = [k for k in calls if k in work]
        dom = max(cand, key=lambda k: calls[k]["ms"] * calls[k]["calls"]) if cand else None
    if dom is None:
        return {"bound": wl.bound, "error": "no library call with a work model was logged", "calls": calls}
    dms = calls[dom]["ms"]
    if wl.bound == "tensor":
        peak, src = bf16_peak()
        ach = work[dom] / dms / 1e9
    elif wl.bound == "hbm":
        peak, src = hbm_peak()
        ach = work[dom] / dms / 1e6
    else:
        peak, src = 148 * 16 * 1.965, "16 MUFU ops/clk/SM x 148 SMs x 1.965 GHz (nominal)"
        ach = work[dom] / dms / 1e6
    t = wl.ncu_traffic.get(dom) if world == 1 else None
    return {"bound": wl.bound, "kernel": dom, "achieved": round(ach, 2), "peak": round(peak, 1), "unit": unit,
            "frac": round(ach / peak, 4), "traffic": t[0] if t else None, "traffic_source": t[1] if t else None,
            "peak_source": src, "algorithmic_work_per_launch": work[dom], "ms_per_launch": dms, "calls": calls}


def free_memory():
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    try:
        torch._C._host_emptyCache()          # pinned host blocks of the previous workload's e2e leg
    except Exception:
        pass


def measure(ctx, wl):
    from vaekit import _lib as L
    args, world, dev = ctx.args, ctx.world, ctx.dev
    warm = max(args.warmup, 3)
    model = wl.model(dev)
    if isinstance(model, torch.nn.Module):
        model = model.to(dev)
    t = wl.make(dev, ctx.rank, world)
    torch.manual_seed(1265)
    for _ in range(warm):
        wl.step(model, t, world)
    sampler = ClockSampler(ctx.local).start() if ctx.rank == 0 else None
    ms, _, launches, d = timed(ctx, lambda: wl.step(model, t, world), args.steps)
    # the same K steps again with CUDA events around every library call (launching stream): per-call durations for the
    # roofline.  Kept out of the pass that gives the step time: two event records per call cost the host 20-60 us.
    out2 = timed(ctx, lambda: wl.step(model, t, world), args.steps, log_events=True)
    log = out2[1]
    loss = {k: float(v.item()) for k, v in d.items()}
    # free the eager autograd graphs (every reference to a returned loss): their AccumulateGrad nodes are bound to this
    # stream and a live one makes the engine synchronise it with the capturing stream, which invalidates the capture
    del d, out2
    import gc
    gc.collect()
    res = {"ms_per_step": round(ms / args.steps, 4), "api": "eager: forward + backward through the drop-in model API"}
    step_ms = ms / args.steps
    gs = None
    if wl.graphed and world == 1:
        # host-bound short steps: the same step captured once in a CUDA graph (vaekit.graphs.GraphedStep) and replayed
        from vaekit.graphs import GraphedStep
        gs = GraphedStep(lambda inp: wl.step(model, inp, world), t, rng=isinstance(wl, VanillaB64))
        for _ in range(warm):
            gs()
        gms, _, _, _ = timed(ctx, gs, args.steps)
        res.update({"eager_ms_per_step": res["ms_per_step"], "graph_ms_per_step": round(gms / args.steps, 4),
                    "ms_per_step": round(gms / args.steps, 4), "vaekit_launches_per_step": launches / args.steps,
                    "api": "vaekit.graphs.GraphedStep: forward + backward through the drop-in model API, captured once, replayed"})
        step_ms = gms / args.steps
    clocks = sampler.stop() if sampler else None
    value = wl.units(world) * 1e3 / step_ms
    roof = roofline(wl, world, log)
    # end to end: pinned host inputs copied in and every returned loss read back, every step
    e2e = None
    if not args.no_e2e:
        host = {k: v.detach().cpu().pin_memory() for k, v in t.items()}     # detached, like a data loader's batch
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        if gs is not None:
            def e2e_step():
                out = gs(**host)
                return [v.item() for v in out.values()]
        else:
            # the copy of batch t+1 runs on a side stream while step t computes (vaekit.io.DevicePrefetcher: the pinned-memory
            # DataLoader of dataset.py:172-178 with its copy taken off the critical path); every timed step still issues
            # one full host -> device copy of its inputs and reads every returned loss back
            from vaekit.io import DevicePrefetcher
            pf = DevicePrefetcher(dev)
            pf.submit(host)

            def e2e_step():
                td = pf.get()
                dd = wl.step(model, td, world)                      # enqueues the step (asynchronous)
                # the next batch's copy goes in AFTER the step's own launches: a step may start with a small host -> device
                # copy of its own (SWAE's projection matrix is drawn on the CPU like the reference's, swae.py:135-149), and
                # that would queue behind a 537 MB prefetch in the copy engine (scripts/probe_prefetch*.py, r02z)
                pf.submit(host)
                return [v.item() for v in dd.values()]              # the caller's .item() per dict key (experiment.py:55)
        for _ in range(2):
            e2e_step()
        ems, _, _, vals = timed(ctx, e2e_step, args.steps)
        e2e = {"value": round(wl.units(world) * args.steps / (ems / 1e3), 1), "unit": "samples/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 * len(vals), "ms_per_step": round(ems / args.steps, 3)}
        del host
    extras = {}
    if ctx.rank == 0 and world == 1 and not args.no_extras:
        extras = wl.extras(model, t, args, dev)
    del t, model, gs
    free_memory()
    cpu = None
    if ctx.rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(wl)
    res.update({"metric": METRIC, "value": round(value, 1), "unit": "samples/s", "scaling": wl.scaling,
                "config": {"workload": wl.desc, "global_batch": wl.units(world), "l2_policy": wl.l2_policy,
                           "parallelism": ("%s; %d ranks" % (wl.parallelism, world)) if world > 1 else "single GPU",
                           "timing": "step: K steps between two CUDA events; roofline: the same K steps repeated with CUDA "
                                     "events around every library call",
                           "e2e_pipeline": "vaekit.io.DevicePrefetcher: batch t+1 copied from pinned memory on a side stream "
                                           "during step t; every step copies its inputs and reads its losses back",
                           "numa_cpus_bound": ctx.numa},
                "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "loss": loss})
    res.update(extras)
    return res


def run_gpu_arm(args):
    ctx = Ctx(args)
    import vaekit
    vaekit.lib()                        # raises if libvaekit.so is missing: there is no fallback
    if args.workload == "all":
        names = ORDER_1GPU if ctx.world == 1 else ORDER_NGPU
    else:
        names = [args.workload]
    results = {}
    for name in names:
        wl = WORKLOADS[name]()
        if ctx.world > 1 or name == names[0]:
            results[name] = measure(ctx, wl)            # a failure of the headline (or on any rank of a group) is fatal
        else:
            try:
                results[name] = measure(ctx, wl)
            except Exception as e:                      # keep the line: one broken workload must not hide the others
                results[name] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
                free_memory()
    if ctx.rank == 0:
        top = results[names[0]]
        line = {"metric": METRIC, "value": top["value"], "unit": "samples/s", "n_gpus": ctx.world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": top["ms_per_step"], "higher_is_better": True,
                "scaling": top["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
        line.update({k: v for k, v in top.items() if k not in line})
        line["headline_workload"] = names[0]
        line["workloads"] = {n: results[n] for n in names[1:]}
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="all", choices=["all"] + sorted(WORKLOADS))
    ap.add_argument("--impl", default="vaekit", choices=["vaekit", "reference"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
