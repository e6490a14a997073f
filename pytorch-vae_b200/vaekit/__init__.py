"""vaekit: B200-native (sm_100a) kernels for the stochastic-latent + loss hot
path of bplaut/PyTorch-VAE, behind the reference's own model-method surface.

    from vaekit.models import vae_models          # same registry keys as models/__init__.py:70-91
    model = vae_models['VanillaVAE'](in_channels=3, latent_dim=128).cuda()
    out = model(x); loss = model.loss_function(*out, M_N=0.005)

Everything that computes runs in libvaekit.so (include/vaekit.h); there is no CPU path.
"""
from . import functional  # noqa: F401
from ._lib import LIB_PATH, declared_symbols, lib  # noqa: F401

__version__ = "0.1.0"
