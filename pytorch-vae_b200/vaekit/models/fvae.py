"""Drop-in for models/fvae.py: FactorVAE.  The VAE branch's reparameterize / KL / reconstruction
are K1 / K2; the MLP discriminator and its cross-entropy branch stay torch (fvae.py:79-89, 182-197)."""
from typing import List

import torch
from torch import Tensor, nn
from torch.nn import functional as TF

from .. import functional as F
from .base import GaussianConvVAE


class FactorVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, gamma: float = 40., **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.gamma = gamma
        self.discriminator = nn.Sequential(nn.Linear(self.latent_dim, 1000), nn.BatchNorm1d(1000), nn.LeakyReLU(0.2),
                                           nn.Linear(1000, 1000), nn.BatchNorm1d(1000), nn.LeakyReLU(0.2),
                                           nn.Linear(1000, 1000), nn.BatchNorm1d(1000), nn.LeakyReLU(0.2),
                                           nn.Linear(1000, 2))
        self.D_z_reserve = None

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var, z]

    def permute_latent(self, z: Tensor) -> Tensor:
        """fvae.py:136-146: an independent permutation of the latent dims of every sample (CPU generator)."""
        B, D = z.size()
        inds = torch.cat([(D * i) + torch.randperm(D) for i in range(B)])
        return z.view(-1)[inds].view(B, D)

    def loss_function(self, *args, **kwargs) -> dict:
        """fvae.py:148-201."""
        recons, input, mu, log_var, z = args[0], args[1], args[2], args[3], args[4]
        kld_weight = kwargs['M_N']
        optimizer_idx = kwargs['optimizer_idx']
        if optimizer_idx == 0:                                   # update the VAE
            elbo, recons_loss, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var),
                                                   combine_mode=F.COMBINE_LINEAR, w=kld_weight)
            self.D_z_reserve = self.discriminator(z)
            vae_tc_loss = (self.D_z_reserve[:, 0] - self.D_z_reserve[:, 1]).mean()
            loss = elbo + self.gamma * vae_tc_loss
            return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'KLD': neg_kld, 'VAE_TC_Loss': vae_tc_loss}
        elif optimizer_idx == 1:                                 # update the discriminator
            device = input.device
            true_labels = torch.ones(input.size(0), dtype=torch.long, requires_grad=False).to(device)
            false_labels = torch.zeros(input.size(0), dtype=torch.long, requires_grad=False).to(device)
            z = z.detach()
            D_z_perm = self.discriminator(self.permute_latent(z))
            D_tc_loss = 0.5 * (TF.cross_entropy(self.D_z_reserve, false_labels) + TF.cross_entropy(D_z_perm, true_labels))
            return {'loss': D_tc_loss, 'D_TC_Loss': D_tc_loss}
