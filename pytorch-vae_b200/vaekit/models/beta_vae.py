"""Drop-in for models/beta_vae.py (loss types 'H' and 'B')."""
from typing import List

import torch
from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE


class BetaVAE(GaussianConvVAE):
    num_iter = 0            # class attribute shadowed by an instance attribute on first call (beta_vae.py:10,132)

    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, beta: int = 4,
                 gamma: float = 1000., max_capacity: int = 25, Capacity_max_iter: int = 1e5,
                 loss_type: str = 'B', **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.beta = beta
        self.gamma = gamma
        self.loss_type = loss_type
        self.C_max = torch.Tensor([max_capacity])
        self.C_stop_iter = Capacity_max_iter

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """beta_vae.py:129-152.  The capacity C is a host scalar (it depends only on the Python
        iteration counter); it is evaluated in fp32 exactly like the reference's one-element tensor."""
        self.num_iter += 1
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        kld_weight = kwargs['M_N']
        kl = self._kl_rows(mu, log_var)
        if self.loss_type == 'H':
            loss, recon, kld, _ = F.elbo(recons, input, kl, combine_mode=F.COMBINE_LINEAR,
                                         w=self.beta * kld_weight)
        elif self.loss_type == 'B':
            C = torch.clamp(self.C_max / self.C_stop_iter * self.num_iter, 0, self.C_max.data[0])   # CPU, [1]
            loss, recon, kld, _ = F.elbo(recons, input, kl, combine_mode=F.COMBINE_ABS,
                                         w=self.gamma * kld_weight, C=float(C[0]))
            loss = loss.view(1)          # the reference's loss inherits C's shape [1]
        else:
            raise ValueError('Undefined loss type.')
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': kld}
