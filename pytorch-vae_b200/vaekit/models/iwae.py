"""Drop-in for models/iwae.py."""
from typing import List

import torch
from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE


class IWAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, num_samples: int = 5,
                 **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.num_samples = num_samples

    def decode(self, z: Tensor) -> Tensor:
        """iwae.py:95-110: [B,S,D] -> [B,S,C,H,W].  z from reparameterize() is dense, so the
        reference's view() (which fails on its own permuted z) is valid here."""
        B = z.size(0)
        r = super().decode(z.reshape(-1, self.latent_dim))
        return r.view([B, -1, r.size(1), r.size(2), r.size(3)])

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        # the reference materialises repeat().permute() (iwae.py:123-124); a stride-0 view carries
        # the same values and the kernel reads it through its row strides
        mu = mu.unsqueeze(1).expand(-1, self.num_samples, -1)
        log_var = log_var.unsqueeze(1).expand(-1, self.num_samples, -1)
        z = self.reparameterize(mu, log_var)
        eps = (z - mu) / log_var            # iwae.py:126 (divides by log_var; unused by the loss)
        return [self.decode(z), input, mu, log_var, z, eps]

    def loss_function(self, *args, **kwargs) -> dict:
        """iwae.py:129-160 as one kernel (vk_iw_bound_fwd): the input image is never repeated."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        kld_weight = kwargs['M_N']
        loss, recon, _, neg_kld = F.iw_bound(recons, input, self._kl_rows(mu, log_var), kld_weight)
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': neg_kld}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        """iwae.py:162-180: one sample dim of size 1 through the [B,S,D] decode, squeezed away again."""
        z = torch.randn(num_samples, 1, self.latent_dim)      # CPU generator, as the reference
        return self.decode(z.to(current_device)).squeeze()

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0][:, 0, :]
