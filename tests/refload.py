"""Import the *reference* model zoo read-only from /root/reference (build
container only; the GPU box has no such path).  Used to pin the oracle and to
generate tests/golden/*.npz.  Never used by -m gpu tests, smoke() or bench.py."""
import contextlib
import os
import sys

import torch

REF_ROOT = "/root/reference"


def have_reference() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "models"))


def load_reference_models():
    sys.dont_write_bytecode = True          # the mount is read-only
    import warnings
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import models as ref_models         # noqa: E402
    return ref_models


@contextlib.contextmanager
def injected_noise(*tensors):
    """Make the reference's ``torch.randn_like`` calls return the given tensors,
    in call order (vanilla_vae.py:116, wae_mmd.py:195, swae.py:165 look the
    function up on the torch module at call time)."""
    queue = list(tensors)
    orig = torch.randn_like

    def fake(t, **kw):
        nxt = queue.pop(0)
        assert nxt.shape == t.shape, (nxt.shape, t.shape)
        return nxt.to(t.dtype)

    torch.randn_like = fake
    try:
        yield
    finally:
        torch.randn_like = orig
