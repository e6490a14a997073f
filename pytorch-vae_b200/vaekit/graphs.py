"""CUDA-graph capture of a whole hot-path step (forward + backward through the vaekit ops).

The short steps are bound by the host, not by the kernels: config 1 (batch 64) moves 16 MB per step (2.5 us of HBM time,
SURVEY 0.10) and the VectorQuantizer step is two kernels of ~40 us each issued by ~180 us of Python.  Every vaekit entry
point only enqueues work on the current stream (no allocation inside the library, no synchronisation, workspace per
stream), so a step is capturable as it stands; the one thing a graph would freeze is the noise, which is why K1 and the
prior draws can read their Philox (seed, offset) from device memory (functional.device_rng).
"""
import torch

from . import functional as F


class GraphedStep:
    """Captures `fn(inputs)` once and replays it.

    fn       callable taking the dict of static input tensors and returning a tensor / tuple / dict of tensors; it
             may call backward() (gradients then live in static graph memory: read `tensor.grad` after a replay)
    inputs   dict of CUDA tensors that stay at fixed addresses; __call__(**new) copies new values into them
    rng      True: the step draws noise -> an int64 device key is made from torch's generator and every replay
             advances it on the device (F.device_rng)
    """

    def __init__(self, fn, inputs, warmup=3, rng=False):
        self.inputs = inputs
        dev = next(iter(inputs.values())).device
        self.key = F.make_rng_key(dev) if rng else None
        self.stream = torch.cuda.Stream(device=dev)
        self.stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self.stream):
            for _ in range(warmup):                       # allocates the per-stream workspace, autotunes cuDNN, ...
                self._run(fn)
        torch.cuda.current_stream(dev).wait_stream(self.stream)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self.outputs = self._run(fn)

    def _run(self, fn):
        if self.key is None:
            return fn(self.inputs)
        with F.device_rng(self.key):
            return fn(self.inputs)

    def __call__(self, **new_inputs):
        with torch.no_grad():                               # the static inputs may be leaves that require grad
            for k, v in new_inputs.items():
                self.inputs[k].copy_(v, non_blocking=True)
        self.graph.replay()
        return self.outputs
