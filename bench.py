#!/usr/bin/env python
"""Benchmark of the latent + loss hot path (BASELINE.json metric: samples/s fwd+bwd).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload iwae|miwae|vanilla] [--impl reference]

Default workload = BASELINE.json configs[1]: IWAE K=5, batch 4096, latent 128, 64x64 RGB --
reparameterize + importance-weighted bound, forward + backward, on synthetic tensors.
One JSON line on stdout (rank 0).  Under torchrun each rank owns its own batch of 4096
(weak scaling, no data-path collective: the op is batch-local).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "pytorch-vae_b200"),):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

P_IMG = 3 * 64 * 64
WORKLOADS = {
    # name: (B, G, K, D, M_N)
    "iwae": dict(B=4096, G=1, K=5, D=128, M_N=0.00025, desc="IWAE K=5, batch 4096, latent 128, 3x64x64 (BASELINE configs[1])"),
    "miwae": dict(B=4096, G=3, K=5, D=128, M_N=0.00025, desc="MIWAE K=5 M=3, batch 4096, latent 128, 3x64x64 (BASELINE configs[1])"),
    "vanilla": dict(B=65536, G=0, K=0, D=128, M_N=0.005, desc="VanillaVAE loss, batch 65536, latent 128, 3x64x64 (config 1 at streaming size)"),
}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


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
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ---------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------
def sample_shape(wl, B):
    if wl["K"] == 0:
        return (B,)
    return (B, wl["G"], wl["K"]) if wl["G"] > 1 else (B, wl["K"])


def make_host_inputs(wl, B, seed, pin):
    g = torch.Generator().manual_seed(seed)
    ss = sample_shape(wl, B)
    t = {
        "recons": torch.tanh(torch.randn(*ss, 3, 64, 64, generator=g)),       # decoder ends in Tanh
        "input": torch.rand(B, 3, 64, 64, generator=g),                       # ToTensor range
        "mu0": torch.randn(B, wl["D"], generator=g),
        "lv0": 0.5 * torch.randn(B, wl["D"], generator=g) - 1.0,
        "gz": 1e-3 * torch.randn(*ss, wl["D"], generator=g),                  # upstream grad of z (decoder bwd)
    }
    if pin:
        t = {k: v.pin_memory() for k, v in t.items()}
    return t


def make_device_inputs(wl, B, seed, dev):
    """Synthetic tensors generated on the device (fast for 1-3 GB), same distributions."""
    g = torch.Generator(device=dev).manual_seed(seed)
    ss = sample_shape(wl, B)
    return {
        "recons": torch.tanh(torch.randn(*ss, 3, 64, 64, device=dev, generator=g)),
        "input": torch.rand(B, 3, 64, 64, device=dev, generator=g),
        "mu0": torch.randn(B, wl["D"], device=dev, generator=g),
        "lv0": 0.5 * torch.randn(B, wl["D"], device=dev, generator=g) - 1.0,
        "gz": 1e-3 * torch.randn(*ss, wl["D"], device=dev, generator=g),
    }


def expand_latents(wl, mu0, lv0):
    if wl["K"] == 0:
        return mu0, lv0
    if wl["G"] > 1:
        shp = (-1, wl["G"], wl["K"], -1)
        return mu0.view(mu0.size(0), 1, 1, -1).expand(*shp), lv0.view(lv0.size(0), 1, 1, -1).expand(*shp)
    return mu0.unsqueeze(1).expand(-1, wl["K"], -1), lv0.unsqueeze(1).expand(-1, wl["K"], -1)


def hot_path_step(model, wl, t):
    """One pass of the hot path through the public model API: reparameterize -> loss_function -> backward."""
    recons = t["recons"].requires_grad_(True)
    mu0 = t["mu0"].requires_grad_(True)
    lv0 = t["lv0"].requires_grad_(True)
    recons.grad = mu0.grad = lv0.grad = None
    mu, lv = expand_latents(wl, mu0, lv0)
    z = model.reparameterize(mu, lv)
    if wl["K"] == 0:
        d = model.loss_function(recons, t["input"], mu, lv, M_N=wl["M_N"])
    else:
        d = model.loss_function(recons, t["input"], mu, lv, z, None, M_N=wl["M_N"])
    torch.autograd.backward([d["loss"], z], [None, t["gz"]])
    return d


def algorithmic_bytes(wl, B):
    """bytes one step must move (fp32)."""
    D = wl["D"]
    if wl["K"] == 0:
        return B * (5 * P_IMG * 4 + 8 * D * 4)
    GK = max(wl["G"], 1) * wl["K"]
    recon = B * ((GK * P_IMG + P_IMG) * 4 * 2 + GK * P_IMG * 4)          # fwd read, bwd read + write g_recons
    lat = B * (2 * D * 4 * 2 + GK * D * 4 * 4)                           # mu0/lv0 read twice; z, g_z, g_mu, g_lv rows
    return recon + lat


def build_model(name):
    from vaekit.models import vae_models
    if name == "iwae":
        return vae_models["IWAE"](3, 128, num_samples=5)
    if name == "miwae":
        return vae_models["MIWAE"](3, 128, num_samples=5, num_estimates=3)
    return vae_models["VanillaVAE"](3, 128)


KERNEL_BYTES = {
    # kernel -> algorithmic bytes per launch as a function of the workload (fp32), SURVEY.md 8(d) / DESIGN.md
    "iw_bound_fwd": lambda B, GK, D: B * (GK * P_IMG + P_IMG) * 4 + B * GK * 8,
    "iw_bound_bwd": lambda B, GK, D: B * (2 * GK * P_IMG + P_IMG) * 4 + B * GK * 12,
    "elbo_fwd": lambda B, GK, D: B * 2 * P_IMG * 4 + B * 4,
    "elbo_bwd": lambda B, GK, D: B * 3 * P_IMG * 4 + B * 4,
    "reparam_kl_fwd": lambda B, GK, D: B * 2 * D * 4 + B * max(GK, 1) * (D * 4 + 4),
    "reparam_kl_bwd": lambda B, GK, D: B * 2 * D * 4 + B * max(GK, 1) * (3 * D * 4 + 4),
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch from `ncu --set full` captures of the SAME workload size
# (profiles/): (kernel, B, G*K) -> bytes.  Anything not captured reports traffic = null.
NCU_TRAFFIC = {
    ("iw_bound_bwd", 4096, 5): (1.208158e9 + 952.969216e6, "profiles/ncu_iw_bound_r01a_summary.csv"),
    ("iw_bound_fwd", 4096, 5): (1.208144e9 + 4.415744e6, "profiles/ncu_iw_bound_r01a_summary.csv"),
}


def roofline_from_events(wl, B, log):
    """Per-kernel average duration from the CUDA events recorded around every libvaekit launch
    inside the timed region (same stream), turned into achieved GB/s on algorithmic bytes."""
    peak, peak_src = peaks()
    GK = max(wl["G"], 1) * wl["K"]
    acc = {}
    for name, e0, e1 in log:
        acc.setdefault(name, []).append(e0.elapsed_time(e1))
    res, total = {}, 0.0
    for name, ts in acc.items():
        ms = sum(ts) / len(ts)
        nbytes = KERNEL_BYTES[name](B, GK, wl["D"]) if name in KERNEL_BYTES else 0
        res[name] = {"ms": round(ms, 4), "launches": len(ts), "bytes": nbytes, "GBps": round(nbytes / ms / 1e6, 1)}
        total += ms
    dom = max(res, key=lambda k: res[k]["ms"])
    for k in res:
        res[k]["share_of_kernel_time"] = round(res[k]["ms"] / total, 4)
    d = res[dom]
    traffic = NCU_TRAFFIC.get((dom, B, GK))
    return {"bound": "hbm", "kernel": dom + "_kernel", "achieved": d["GBps"], "peak": peak, "unit": "GB/s",
            "frac": round(d["GBps"] / peak, 4), "traffic": traffic[0] if traffic else None,
            "traffic_source": traffic[1] if traffic else None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": d["bytes"], "ms_per_launch": d["ms"], "kernels": res}


# ---------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference's torch code, on the host cores
# ---------------------------------------------------------------------------
def cpu_reference_step(wl, t):
    """reparameterize + loss_function + backward exactly as the reference composes them
    (iwae.py:121-160 / miwae.py:124-164 / vanilla_vae.py:119-146), via oracle/vae_oracle.py."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import vae_oracle as O
    recons = t["recons"].detach().requires_grad_(True)
    mu0 = t["mu0"].detach().requires_grad_(True)
    lv0 = t["lv0"].detach().requires_grad_(True)
    if wl["K"] == 0:
        mu, lv = mu0, lv0
    elif wl["G"] > 1:
        mu = mu0.repeat(wl["G"], wl["K"], 1, 1).permute(2, 0, 1, 3)
        lv = lv0.repeat(wl["G"], wl["K"], 1, 1).permute(2, 0, 1, 3)
    else:
        mu = mu0.repeat(wl["K"], 1, 1).permute(1, 0, 2)
        lv = lv0.repeat(wl["K"], 1, 1).permute(1, 0, 2)
    z = O.reparameterize(mu, lv, torch.randn_like(mu))
    if wl["K"] == 0:
        d = O.vanilla_loss(recons, t["input"], mu, lv, wl["M_N"])
    else:
        d = O.iw_bound_loss(recons, t["input"], mu, lv, wl["M_N"])
    torch.autograd.backward([d["loss"], z], [None, t["gz"]])
    return d


def cpu_baseline(wl, sample_B, iters, warmup=1):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    t = make_host_inputs(wl, sample_B, 1265, pin=False)
    for _ in range(warmup):
        cpu_reference_step(wl, t)
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        cpu_reference_step(wl, t)
        ts.append(time.perf_counter() - t0)
    return ts, cores


def run_reference_arm(args, wl):
    """--impl reference: the reference's CPU path (oracle port of its torch code; the Python reference
    itself cannot travel to the GPU box) on all host cores, bounded sample of the same workload."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    sample_B = args.cpu_sample or (1024 if wl["K"] else 4096)
    ts, cores = cpu_baseline(wl, sample_B, args.steps, warmup=max(1, min(args.warmup, 2)))
    total = sum(ts)
    value = sample_B * len(ts) / total
    line = {
        "impl": "reference", "metric": "latent+loss samples/s fwd+bwd", "value": round(value, 2), "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * total / len(ts), 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "sample": "batch %d of %d per step" % (sample_B, wl["B"])},
        "cpu_baseline": {"value": round(value, 2), "unit": "samples/s", "cores": cores, "kind": "port",
                         "sample": "batch %d of the %d-sample workload per step, torch CPU %d threads" % (sample_B, wl["B"], cores)},
        "e2e": {"value": round(value, 2), "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def run_gpu_arm(args, wl):
    import torch.distributed as dist
    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import vaekit
    from vaekit import _lib as L
    vaekit.lib()
    model = build_model(args.workload)
    B = wl["B"]
    t = make_device_inputs(wl, B, 1265 + rank, dev)
    torch.manual_seed(1265)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident timing (value) ----
    for _ in range(args.warmup):
        hot_path_step(model, wl, t)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = L.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        d = hot_path_step(model, wl, t)
    e1.record()
    barrier()
    launches = L.LAUNCHES - launches0
    ms = e0.elapsed_time(e1)
    # the same K steps again with CUDA events around every library call (launching stream): per-kernel durations for the
    # roofline.  Kept out of the pass that gives `value`: two event records per call cost the host 20-60 us per call.
    L.EVENT_LOG = []
    for _ in range(args.steps):
        hot_path_step(model, wl, t)
    barrier()
    log, L.EVENT_LOG = L.EVENT_LOG, None
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        tms = torch.tensor([ms], device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = tms.item()
    value = world * B * args.steps / (ms / 1e3)
    roof = roofline_from_events(wl, B, log)
    loss_val = d["loss"].item()

    # ---- end-to-end: pinned host tensors in, loss out, every step ----
    host = make_host_inputs(wl, B, 99 + rank, pin=True) if not args.no_e2e else None
    e2e = None
    if host is not None:
        h2d = sum(v.numel() * 4 for v in host.values())

        def e2e_step():
            td = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
            dd = hot_path_step(model, wl, td)
            return [dd[k].item() for k in dd]          # the caller's .item() per dict key (experiment.py:55)

        for _ in range(min(args.warmup, 3)):
            e2e_step()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(args.steps):
            vals = e2e_step()
        f1.record()
        barrier()
        ems = f0.elapsed_time(f1)
        if world > 1:
            tms = torch.tensor([ems], device=dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ems = tms.item()
        e2e = {"value": round(world * B * args.steps / (ems / 1e3), 1), "unit": "samples/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 * len(vals), "ms_per_step": round(ems / args.steps, 3)}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            sample_B = args.cpu_sample or (1024 if wl["K"] else 4096)
            ts, cores = cpu_baseline(wl, sample_B, 3)
            v = sample_B * len(ts) / sum(ts)
            cpu = {"value": round(v, 2), "unit": "samples/s", "cores": cores, "kind": "port",
                   "sample": "batch %d of the %d-sample workload, 3 iterations after 1 warm-up, torch CPU %d threads"
                             % (sample_B, B, cores)}
        alg = algorithmic_bytes(wl, B)
        line = {
            "metric": "latent+loss samples/s fwd+bwd", "value": round(value, 1), "unit": "samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "per_gpu_batch": B, "global_batch": B * world,
                       "l2_policy": "inputs (%.2f GB/step) larger than L2, no flush" % (alg / 1e9),
                       "parallelism": "batch-sharded, no data-path collective" if world > 1 else "single GPU",
                       "step": "reparameterize + loss_function + backward through the drop-in model API",
                       "timing": "value: K steps between two CUDA events; roofline: the same K steps repeated with CUDA "
                                 "events around every library call"},
            "step_GBps_algorithmic": round(alg / (ms / args.steps) / 1e6, 1),
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "loss": loss_val,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------
# op workloads: the other BASELINE.json configs (WAE-MMD, VQ, SWAE) and BetaTC, through the drop-in model API
# ---------------------------------------------------------------------------
def bf16_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["bf16_tflops_sustained"]), "MEASURED_PEAKS.json bf16_tflops_sustained (measured)"
    except Exception:
        return 1400.0, "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)"


class OpWorkload:
    """One hot-path op at a BASELINE.json size.  make() builds synthetic inputs, step() runs forward + backward through
    the public model API, cpu_step() is the oracle port of the reference on a bounded sample."""
    scaling = "weak"

    def units(self, world):
        return self.N * (world if self.scaling == "weak" else 1)


class WaeMmd(OpWorkload):
    scaling = "strong"          # fixed global batch, rows sharded after an all-gather of z

    def __init__(self, kt):
        self.kt, self.N, self.D = kt, 65536, 128
        self.reg = 5000.0 if kt == "rbf" else 100.0
        self.desc = "WAE-MMD %s kernel, z batch 65536x128, value + d/dz in one tcgen05 Gram sweep (BASELINE configs[2])" % kt.upper()
        self.cpu_sample, self.bound = 2048, "tensor"
        self.dominant = "mmd_fwd"

    def model(self):
        from vaekit.models import vae_models
        return vae_models["WAE_MMD"](3, self.D, reg_weight=self.reg, kernel_type=self.kt)

    def make(self, dev, rank, world, n=None, generator_device=None):
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
        return {"MMD": mmd}

    def work(self, world):
        return {"mmd_fwd": 10.0 * self.N * self.N * self.D / world}, "TFLOP/s"

    def cpu_step(self, t):
        import vae_oracle as O
        z = t["z"].detach().requires_grad_(True)
        n = z.shape[0]
        O.mmd(z, torch.randn_like(z), self.reg / (n * (n - 1)), self.kt, 2.0).backward()


class VqWorkload(OpWorkload):
    def __init__(self):
        self.N, self.D, self.K, self.H = 1024, 64, 512, 16
        self.desc = "VQ-VAE VectorQuantizer, codebook 512x64, 16x16 latent grid, batch 1024 (BASELINE configs[3])"
        self.cpu_sample, self.bound = 128, "hbm"
        self.dominant = "vq_fwd"

    def model(self):
        from vaekit.models.vq_vae import VectorQuantizer
        m = VectorQuantizer(self.K, self.D, 0.25)
        with torch.no_grad():
            m.embedding.weight.normal_()
        return m

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
        return {"VQ_Loss": vq_loss}

    def work(self, world):
        R = self.N * self.H * self.H
        fwd = 2 * R * self.D * 4 + R * 4 + self.K * self.D * 4
        bwd = 4 * R * self.D * 4 + R * 4 + 2 * self.K * self.D * 4
        return {"vq_fwd": float(fwd), "vq_bwd": float(bwd)}, "GB/s"

    def cpu_step(self, t):
        import vae_oracle as O
        x = t["latents"].detach().requires_grad_(True)
        cb = torch.randn(self.K, self.D, requires_grad=True)
        q, l, _ = O.vq_forward(x, cb, 0.25)
        torch.autograd.backward([q, l], [t["gq"], None])


class SwaeWorkload(OpWorkload):
    scaling = "strong"          # fixed global batch, projections sharded across ranks

    def __init__(self):
        self.N, self.D, self.S = 1 << 20, 128, 200
        self.desc = "SWAE sliced-Wasserstein, 200 projections, batch 1048576x128, projection GEMM + radix sort (BASELINE configs[4])"
        self.cpu_sample, self.bound = 1 << 16, "hbm"
        self.dominant, self.dominant_sharded = "swd_fwd", "swd_segments_fwd"
        # dram__bytes_read.sum + dram__bytes_write.sum over the 19 kernels of one vk_swd_fwd call at this size
        # (`ncu --set full`, r01v): 30.82 GB against the 28.76 GB of the traffic model
        self.ncu_traffic = {"swd_fwd": (30.822819e9, "profiles/ncu_swae_r01v_summary.txt")}

    def model(self):
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
        return {"SWD": swd}

    def work(self, world):
        # radix-sort traffic model of SURVEY 8(d); sharded (all-to-all) runs time the sort + loss stage on its own:
        # 120 of the model's 132 bytes per (sample, projection) belong to it (key+index sort 68, key-only sort 36,
        # loss read 8, gradient scatter 8)
        return {"swd_fwd": 27424.0 * self.N / world,
                "swd_segments_fwd": 120.0 * self.N * self.S / world}, "GB/s"

    def cpu_step(self, t):
        import vae_oracle as O
        z = t["z"].detach().requires_grad_(True)
        n = z.shape[0]
        O.swd(z, torch.randn_like(z), O.random_projections(self.D, self.S), 2.0, 100.0 / (n * (n - 1))).backward()


class BetaTcWorkload(OpWorkload):
    scaling = "strong"          # fixed global batch, anchor rows sharded after an all-gather of z, mu, log_var

    def __init__(self):
        self.N, self.D = 4096, 128
        self.desc = "BetaTC-VAE latent terms (log_q_z, per-dim LSE, MI/TC/KLD), batch 4096, latent 128 (SURVEY 8(d) extra)"
        self.cpu_sample, self.bound = 1024, "mufu"
        self.dominant, self.dominant_sharded = "tc_bwd", "tc_bwd_rows"

    def model(self):
        from vaekit.models import vae_models
        return vae_models["BetaTCVAE"](3, self.D).eval()

    def make(self, dev, rank, world, n=None):
        n = n or self.N // world
        g = torch.Generator(device=dev).manual_seed(1265 + rank)
        mu = torch.randn(n, self.D, device=dev, generator=g)
        lv = 0.3 * torch.randn(n, self.D, device=dev, generator=g)
        return {"mu": mu, "lv": lv, "eps": torch.randn(n, self.D, device=dev, generator=g)}

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
        return {"latent": latent}

    def work(self, world):
        e = float(self.N) * self.N * self.D
        return {"tc_fwd": e, "tc_bwd": 2.0 * e, "tc_fwd_rows": e / world, "tc_bwd_rows": 2.0 * e / world}, "Gexp/s"

    def cpu_step(self, t):
        import vae_oracle as O
        mu, lv = t["mu"].detach().requires_grad_(True), t["lv"].detach().requires_grad_(True)
        z = O.reparameterize(mu, lv, t["eps"])
        rz = torch.zeros(mu.shape[0], 1, 1, 1)
        O.betatc_loss(rz, rz, mu, lv, z, 0.00025, alpha=1.0, beta=6.0, gamma=1.0, anneal_rate=1.0)["loss"].backward()


OP_WORKLOADS = {"wae_mmd_rbf": lambda: WaeMmd("rbf"), "wae_mmd_imq": lambda: WaeMmd("imq"), "vq": VqWorkload,
                "swae": SwaeWorkload, "betatc": BetaTcWorkload}


def op_cpu_baseline(wl, iters, warmup=1):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    t = wl.make(torch.device("cpu"), 0, 1, n=wl.cpu_sample)
    for _ in range(warmup):
        wl.cpu_step(t)
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        wl.cpu_step(t)
        ts.append(time.perf_counter() - t0)
    v = wl.cpu_sample * len(ts) / sum(ts)
    return v, cores, ts


def run_op_reference_arm(args, wl):
    rank, _, _ = dist_env()
    if rank != 0:
        return
    v, cores, ts = op_cpu_baseline(wl, max(1, min(args.steps, 3)), warmup=1)
    sample = "batch %d of the %d-sample workload per step (the reference's broadcast formulation at that size), torch CPU %d threads" % (
        wl.cpu_sample, wl.N, cores)
    line = {"impl": "reference", "metric": "latent+loss samples/s fwd+bwd", "value": round(v, 2), "unit": "samples/s",
            "n_gpus": args.gpus, "steps": len(ts), "warmup": 1, "ms_per_step": round(1e3 * sum(ts) / len(ts), 3),
            "higher_is_better": True, "scaling": wl.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.desc, "sample": sample},
            "cpu_baseline": {"value": round(v, 2), "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(v, 2), "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_op_gpu_arm(args, wl):
    import torch.distributed as dist
    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import vaekit
    from vaekit import _lib as L
    vaekit.lib()
    model = wl.model().to(dev)
    t = wl.make(dev, rank, world)
    torch.manual_seed(1265)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps, log_events):
        barrier()
        if log_events:
            L.EVENT_LOG = []
        n0 = L.LAUNCHES
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        barrier()
        log, L.EVENT_LOG = L.EVENT_LOG, None
        ms = e0.elapsed_time(e1)
        if world > 1:
            tms = torch.tensor([ms], device=dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = tms.item()
        return ms, log, L.LAUNCHES - n0, r

    for _ in range(max(args.warmup, 3)):
        wl.step(model, t, world)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, _, launches, d = timed(lambda: wl.step(model, t, world), args.steps, False)
    # the same K steps again with CUDA events around every library call (launching stream): per-call durations for the
    # roofline.  Kept out of the pass that gives `value`: two event records per call cost the host 20-60 us per call.
    _, log, _, _ = timed(lambda: wl.step(model, t, world), args.steps, True)
    clocks = sampler.stop() if rank == 0 else None
    value = wl.units(world) * args.steps / (ms / 1e3)
    # roofline of the dominant library call, from CUDA events around each call on the launching stream
    work, unit = wl.work(world)
    acc = {}
    for name, a, b in log:
        acc.setdefault(name, []).append(a.elapsed_time(b))
    kern = {k: {"ms": round(sum(v) / len(v), 4), "calls": len(v)} for k, v in acc.items()}
    tot = sum(k["ms"] * k["calls"] for k in kern.values()) or 1.0
    for k in kern:
        kern[k]["share_of_library_time"] = round(kern[k]["ms"] * kern[k]["calls"] / tot, 4)
    dom = wl.dominant if wl.dominant in kern else wl.dominant_sharded
    dms = kern[dom]["ms"]
    if wl.bound == "tensor":
        peak, src = bf16_peak()
        ach = work[dom] / dms / 1e9
    elif wl.bound == "hbm":
        peak, src = peaks()
        ach = work[dom] / dms / 1e6
    else:
        peak, src = 148 * 16 * 1.965, "16 MUFU ops/clk/SM x 148 SMs x 1.965 GHz (nominal)"
        ach = work[dom] / dms / 1e6
    ncu_t = getattr(wl, "ncu_traffic", {}).get(dom) if world == 1 else None
    roof = {"bound": "tensor" if wl.bound == "tensor" else "hbm" if wl.bound == "hbm" else "mufu", "kernel": dom,
            "achieved": round(ach, 2), "peak": round(peak, 1), "unit": unit, "frac": round(ach / peak, 4),
            "traffic": ncu_t[0] if ncu_t else None, "traffic_source": ncu_t[1] if ncu_t else None,
            "peak_source": src, "algorithmic_work_per_launch": work[dom], "ms_per_launch": dms, "calls": kern}
    # end to end: host (pinned) inputs copied in and the loss read back every step
    e2e = None
    if not args.no_e2e:
        # detached: the workload marks its inputs requires_grad in place, and a host copy that stayed in the autograd
        # graph would make every backward() end with a pageable device-to-host copy of the gradient (r01w and earlier
        # ops-workload e2e numbers carry that artefact: 1.7-3.4 GB/s effective instead of the 54 GB/s of pinned copies)
        host = {k: v.detach().cpu().pin_memory() for k, v in t.items()}
        h2d = sum(v.numel() * v.element_size() for v in host.values())

        def e2e_step():
            td = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
            dd = wl.step(model, td, world)
            return [v.item() for v in dd.values()]

        for _ in range(2):
            e2e_step()
        ems, _, _, vals = timed(e2e_step, args.steps, False)
        e2e = {"value": round(wl.units(world) * args.steps / (ems / 1e3), 1), "unit": "samples/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 * len(vals), "ms_per_step": round(ems / args.steps, 3)}
    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            v, cores, ts = op_cpu_baseline(wl, 2)
            cpu = {"value": round(v, 2), "unit": "samples/s", "cores": cores, "kind": "port",
                   "sample": "batch %d of the %d-sample workload (the reference's formulation at that size), 2 iterations "
                             "after 1 warm-up, torch CPU %d threads" % (wl.cpu_sample, wl.N, cores)}
        line = {"metric": "latent+loss samples/s fwd+bwd", "value": round(value, 1), "unit": "samples/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 4),
                "higher_is_better": True, "scaling": wl.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": wl.desc, "global_batch": wl.units(world),
                           "l2_policy": "synthetic inputs regenerated never; working set (operand images / sort buffers) larger than L2 or tensor-bound",
                           "parallelism": ("%s over %d ranks" % ("row-sharded after all-gather" if isinstance(wl, (WaeMmd, BetaTcWorkload)) else
                                                                  "projection-sharded, exchange fused into the GEMMs over peer memory (%s)"
                                                                  % os.environ.get("VAEKIT_SWD_EXCHANGE", "auto") if isinstance(wl, SwaeWorkload)
                                                                  else "batch-sharded", world)) if world > 1 else "single GPU",
                           "step": "forward + backward through the drop-in model API",
                           "timing": "value: K steps between two CUDA events; roofline: the same K steps repeated with CUDA "
                                     "events around every library call"},
                "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
                "loss": {k: v.item() for k, v in d.items()}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="iwae", choices=sorted(WORKLOADS) + sorted(OP_WORKLOADS))
    ap.add_argument("--impl", default="vaekit", choices=["vaekit", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="batch of the bounded CPU-baseline sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.workload in OP_WORKLOADS:
        wl = OP_WORKLOADS[args.workload]()
        (run_op_reference_arm if args.impl == "reference" else run_op_gpu_arm)(args, wl)
        return
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl)
    else:
        run_gpu_arm(args, wl)


if __name__ == "__main__":
    main()
