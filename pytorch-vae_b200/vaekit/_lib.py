"""ctypes binding of libvaekit.so (include/vaekit.h).  There is NO fallback: if the
library is missing or a call fails, a RuntimeError is raised."""
import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VAEKIT_LIB") or os.path.join(_HERE, "_lib", "libvaekit.so")   # override: instrumented builds

_p = C.c_void_p
_i64 = C.c_int64
_u64 = C.c_uint64
_f = C.c_float
_i = C.c_int

_SIGNATURES = {
    "vk_version": ([], _i),
    "vk_last_error_string": ([], C.c_char_p),
    "vk_workspace_bytes": ([], _i64),
    "vk_reparam_kl_fwd": ([_p, _p, _p, _p, _p, _i64, _u64, _u64, _p, _p, _p, _p], _i),
    "vk_reparam_kl_bwd": ([_p, _p, _p, _p, _p, _p, _p, _i64, _u64, _u64, _p, _p, _p], _i),
    "vk_reparam_kl_fwd_dev": ([_p, _p, _p, _p, _p, _i64, _p, _p, _p, _p, _p], _i),
    "vk_reparam_kl_bwd_dev": ([_p, _p, _p, _p, _p, _p, _p, _i64, _p, _p, _p, _p], _i),
    "vk_philox_normal": ([_p, _i64, _u64, _u64, _p], _i),
    "vk_philox_normal_dev": ([_p, _i64, _p, _i64, _p], _i),
    "vk_philox_raw": ([_p, _i64, _u64, _u64, _p], _i),
    "vk_elbo_fwd": ([_p, _p, _i64, _p, _i64, _i, _f, _i, _f, _f, _p, _p, _i64, _p], _i),
    "vk_elbo_bwd": ([_p, _p, _i64, _i64, _i, _f, _i, _f, _f, _p, _p, _p, _p, _p, _p, _p, _p], _i),
    "vk_per_image_mse": ([_p, _p, _i64, _i64, _p, _p], _i),
    "vk_iw_bound_fwd": ([_p, _p, _p, _i64, _i64, _i64, _i64, _f, _p, _p, _p, _i64, _p], _i),
    "vk_iw_bound_fwd_static": ([_p, _p, _p, _i64, _i64, _i64, _i64, _f, _p, _p, _p], _i),
    "vk_vq_workspace_bytes": ([_i64, _i64], _i64),
    "vk_vq_fwd": ([_p, _p, _i64, _i64, _i64, _i64, _f, _p, _p, _p, _p, _i64, _p], _i),
    "vk_vq_bwd": ([_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _f, _p, _p, _p], _i),
    "vk_mmd_workspace_bytes": ([_i64, _i64], _i64),
    "vk_mmd_fwd": ([_p, _p, _i64, _i64, _i, _f, _i, _i64, _i64, _p, _p, _p, _p, _i64, _p], _i),
    "vk_mmd_combine": ([_p, _i64, _i, _f, _p, _p], _i),
    "vk_scale_by_scalar": ([_p, _p, _f, _i64, _p, _p], _i),
    "vk_pair_kernel_fwd": ([_p, _p, _p, _p, _i64, _i64, _i64, _i, _f, _f, _p, _p, _p, _i64, _p], _i),
    "vk_pair_kernel_bwd_rows": ([_p, _i64, _i64, _p, _i64, _i64, _i64, _i64, _i64, _i, _f, _f, _p, _p, _i64, _i64, _p, _p,
                                _p], _i),
    "vk_pair_kernel_bwd_dense": ([_p, _p, _p, _p, _i64, _i64, _i64, _i, _f, _f, _p, _p, _i64, _i64, _p, _p, _p], _i),
    "vk_swd_workspace_bytes": ([_i64, _i64], _i64),
    "vk_swd_fwd": ([_p, _p, _p, _i64, _i64, _i64, _f, _p, _p, _p, _i64, _p], _i),
    "vk_mean_scale": ([_p, C.c_double, _f, _p, _p], _i),
    "vk_swd_segments_workspace_bytes": ([_i64, _i64], _i64),
    "vk_swd_segments_fwd": ([_p, _p, _i64, _i64, _f, _p, _i, _p, _i64, _p], _i),
    "vk_seg_sort_workspace_bytes": ([_i64, _i64], _i64),
    "vk_seg_sort": ([_p, _p, _i64, _i64, _p, _i64, _p], _i),
    "vk_gemm_nt_workspace_bytes": ([], _i64),
    "vk_gemm_nt": ([_p, _i64, _i64, _p, _i64, _i64, _p, _i64, _i64, _i64, _i64, _i64, _p, _i64, _p], _i),
    "vk_gemm_nt_tab": ([_p, _p, _i64, _i64, _p, _i64, _i64, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _i64, _p], _i),
    "vk_tc_workspace_bytes": ([_i64, _i64], _i64),
    "vk_tc_fwd": ([_p, _p, _p, _i64, _i64, _f, _f, _f, _f, _p, _p, _p, _p, _i64, _p], _i),
    "vk_tc_bwd": ([_p, _p, _p, _p, _p, _i64, _i64, _f, _f, _f, _f, _p, _p, _p, _p, _p, _i64, _p], _i),
    "vk_tc_fwd_rows": ([_p, _p, _p, _i64, _i64, _i64, _i64, _f, _f, _f, _f, _p, _p, _p, _p, _i64, _p], _i),
    "vk_tc_bwd_rows": ([_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _f, _f, _f, _f, _p, _p, _p, _p, _p, _i64, _p], _i),
    "vk_gumbel_softmax_fwd": ([_p, _i64, _i64, _f, _f, _u64, _u64, _p, _p, _p, _p], _i),
    "vk_gumbel_softmax_bwd": ([_p, _p, _i64, _i64, _f, _p, _p], _i),
    "vk_cat_kl_rows_fwd": ([_p, _i64, _i64, _i64, _f, _p, _p], _i),
    "vk_cat_kl_rows_bwd": ([_p, _p, _i64, _i64, _i64, _f, _p, _p], _i),
    "vk_gauss_logq_rows_fwd": ([_p, _p, _p, _i64, _i64, _p, _p], _i),
    "vk_gauss_logq_rows_bwd": ([_p, _p, _p, _p, _i64, _i64, _p, _p, _p, _p], _i),
    "vk_vamp_prior_fwd": ([_p, _p, _p, _i64, _i64, _i64, _p, _p, _p], _i),
    "vk_vamp_prior_bwd": ([_p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _p], _i),
    "vk_iw_bound_bwd": ([_p, _p, _p, _p, _i64, _i64, _i64, _i64, _f, _p, _p, _p, _p, _p, _p, _p], _i),
}

_lib = None
_lock = threading.Lock()


def lib():
    """The loaded shared library (loads on first use; raises if it was never built)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        "vaekit: %s not found -- build it with `python pytorch-vae_b200/build.py` "
                        "(there is no CPU or eager fallback)" % LIB_PATH)
                handle = C.CDLL(LIB_PATH)
                for name, (args, res) in _SIGNATURES.items():
                    fn = getattr(handle, name)      # AttributeError if the symbol is missing
                    fn.argtypes = args
                    fn.restype = res
                _lib = handle
    return _lib


def declared_symbols():
    return sorted(_SIGNATURES)


def check(rc, what):
    if rc != 0:
        msg = lib().vk_last_error_string().decode("utf-8", "replace")
        raise RuntimeError("vaekit %s failed (%d): %s" % (what, rc, msg))


EVENT_LOG = None      # set to a list to get (name, start_event, end_event) per kernel launch
LAUNCHES = 0          # number of libvaekit kernel launches made by this process


def call(name, *args):
    """Invoke vk_<name>(*args) on the current device/stream; raises on a non-zero return code."""
    global LAUNCHES
    fn = getattr(lib(), "vk_" + name)
    if EVENT_LOG is None:
        rc = fn(*args)
    else:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*args)
        e1.record()
        EVENT_LOG.append((name, e0, e1))
    LAUNCHES += 1
    check(rc, name)


def ptr(t):
    return None if t is None else t.data_ptr()


def stream_ptr():
    return _raw_stream(torch.cuda.current_device())


# Host fast paths: the public torch.cuda.current_stream() / torch.cuda.device() helpers spend 5-10 us each in Python
# argument handling, which is visible on the short steps (VQ: ~180 us of host time per step).  The underlying C entry
# points are used directly when this torch has them.
if hasattr(torch._C, "_cuda_getCurrentRawStream"):
    def _raw_stream(index):
        return torch._C._cuda_getCurrentRawStream(index)
else:                                                    # pragma: no cover
    def _raw_stream(index):
        return torch.cuda.current_stream(index).cuda_stream


class on:
    """`with on(device):` == `with torch.cuda.device(device):` without the argument parsing."""
    __slots__ = ("idx", "prev")

    def __init__(self, device):
        self.idx = device.index if device.index is not None else -1
        self.prev = -1

    def __enter__(self):
        self.prev = torch.cuda._exchange_device(self.idx)

    def __exit__(self, *exc):
        self.idx = torch.cuda._maybe_exchange_device(self.prev)
        return False


if not (hasattr(torch.cuda, "_exchange_device") and hasattr(torch.cuda, "_maybe_exchange_device")):   # pragma: no cover
    on = torch.cuda.device


def require(t, name, dims=None):
    """Host-side argument policy (SURVEY 8(b) 'Errors'): CUDA fp32 tensors only."""
    if not isinstance(t, torch.Tensor):
        raise RuntimeError("vaekit: %s must be a tensor" % name)
    if not t.is_cuda:
        raise RuntimeError("vaekit: %s must be a CUDA tensor (no CPU fallback), got %s" % (name, t.device))
    if t.dtype != torch.float32:
        raise RuntimeError("vaekit: %s must be float32, got %s" % (name, t.dtype))
    if dims is not None and t.dim() not in dims:
        raise RuntimeError("vaekit: %s must have %s dims, got %d" % (name, dims, t.dim()))
    return t


_workspaces = {}
_ws_min = None
_ws_sizes = {}


def ws_bytes(op, *shape):
    """vk_<op>_workspace_bytes(*shape), cached per shape (a ctypes round trip per call otherwise)."""
    key = (op,) + shape
    v = _ws_sizes.get(key)
    if v is None:
        v = int(getattr(lib(), "vk_%s_workspace_bytes" % op)(*shape))
        _ws_sizes[key] = v
    return v


def workspace(device, min_bytes=0):
    """Zero-initialised scratch buffer, one per (device, stream); kernels leave the
    ticket header zeroed, so it is reused across calls without a memset."""
    global _ws_min
    index = device.index if device.index is not None else torch.cuda.current_device()
    key = (index, _raw_stream(index))
    if _ws_min is None:
        _ws_min = int(lib().vk_workspace_bytes())
    need = max(_ws_min, int(min_bytes))
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(need, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def i64x3(vals):
    return (C.c_int64 * 3)(*[int(v) for v in vals])
