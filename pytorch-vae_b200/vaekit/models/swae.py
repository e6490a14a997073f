"""Drop-in for models/swae.py (sliced-Wasserstein autoencoder)."""
from typing import List

import torch
from torch import Tensor
from torch import distributions as dist

from .. import functional as F
from .base import DeterministicConvAE


class SWAE(DeterministicConvAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, reg_weight: int = 100,
                 wasserstein_deg: float = 2., num_projections: int = 50, projection_dist: str = 'normal',
                 **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.reg_weight = reg_weight
        self.p = wasserstein_deg
        self.num_projections = num_projections
        self.proj_dist = projection_dist

    def loss_function(self, *args, **kwargs) -> dict:
        """swae.py:110-127: mse + l1 (one streaming kernel) + SWD; does not read M_N."""
        recons, input, z = args[0], args[1], args[2]
        batch_size = input.size(0)
        bias_corr = batch_size * (batch_size - 1)
        reg_weight = self.reg_weight / bias_corr
        _, recons_loss, _, _ = F.elbo(recons, input, recon_mode=F.RECON_MSE_L1_MEAN)
        swd_loss = self.compute_swd(z, self.p, reg_weight)
        loss = recons_loss + swd_loss
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'SWD': swd_loss}

    def get_random_projections(self, latent_dim: int, num_samples: int) -> Tensor:
        """swae.py:129-148: [S, D] unit directions drawn on the CPU generator (25 600 numbers at the benchmark
        size); kept as host torch code so torch.manual_seed reproduces the reference's directions bit for bit."""
        if self.proj_dist == 'normal':
            raw = torch.randn(num_samples, latent_dim)
        elif self.proj_dist == 'cauchy':
            raw = dist.Cauchy(torch.tensor([0.0]), torch.tensor([1.0])).sample((num_samples, latent_dim)).squeeze()
        else:
            raise ValueError('Unknown projection distribution.')
        return raw / raw.norm(dim=1).view(-1, 1)

    def compute_swd(self, z: Tensor, p: float, reg_weight: float) -> Tensor:
        """swae.py:151-179: prior_z from the contract Philox stream; projection GEMM, per-projection radix sort,
        loss and d/dz in one library call (vk_swd_fwd)."""
        prior_z = F.philox_normal(z.shape, z.device)
        proj = self.get_random_projections(self.latent_dim, num_samples=self.num_projections)
        # through pinned memory, non-blocking: a pageable .to(device) would synchronise the stream on every call
        proj = proj.contiguous().pin_memory().to(z.device, non_blocking=True) if z.is_cuda else proj.to(z.device)
        self.last_prior_z, self.last_proj = prior_z, proj
        return F.swd(z, prior_z, proj, p, reg_weight)
