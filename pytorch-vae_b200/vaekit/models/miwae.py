"""Drop-in for models/miwae.py."""
from typing import List

import torch
from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE


class MIWAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, num_samples: int = 5,
                 num_estimates: int = 5, **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.num_samples = num_samples
        self.num_estimates = num_estimates

    def decode(self, z: Tensor) -> Tensor:
        """miwae.py:97-112: [B,M,S,D] -> [B,M,S,C,H,W]."""
        B, M, S, D = z.size()
        r = super().decode(z.contiguous().view(-1, self.latent_dim))
        return r.view([B, M, S, r.size(-3), r.size(-2), r.size(-1)])

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        shape = (-1, self.num_estimates, self.num_samples, -1)
        mu = mu.view(mu.size(0), 1, 1, -1).expand(*shape)            # values of miwae.py:126-127
        log_var = log_var.view(log_var.size(0), 1, 1, -1).expand(*shape)
        z = self.reparameterize(mu, log_var)
        eps = (z - mu) / log_var            # miwae.py:129
        return [self.decode(z), input, mu, log_var, z, eps]

    def loss_function(self, *args, **kwargs) -> dict:
        """miwae.py:132-164 (mean over B then M == global mean of the per-(b,m) sums)."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        kld_weight = kwargs['M_N']
        loss, recon, _, neg_kld = F.iw_bound(recons, input, self._kl_rows(mu, log_var), kld_weight)
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': neg_kld}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        """miwae.py:166-183: [N,1,1,D] through the [B,M,S,D] decode, squeezed."""
        z = torch.randn(num_samples, 1, 1, self.latent_dim)   # CPU generator, as the reference
        return self.decode(z.to(current_device)).squeeze()

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0][:, 0, 0, :]
