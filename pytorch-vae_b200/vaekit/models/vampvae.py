"""Drop-in for models/vampvae.py: VAE with a VampPrior.  reparameterize is K1, the reconstruction
term and the loss combine are K2, and the two log-density terms of vampvae.py:136-158 are the K9 kernels (the reference's
[B, K, D] broadcast is never materialised).  The pseudo-input embedding and the conv stacks stay torch modules."""
from typing import List

import torch
from torch import Tensor, nn

from .. import functional as F
from .base import GaussianConvVAE


class VampVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, num_components: int = 50,
                 **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.num_components = num_components
        self.pseudo_input = torch.eye(num_components, requires_grad=False)       # plain attribute, vampvae.py:78
        self.embed_pseudo = nn.Sequential(nn.Linear(num_components, 12288), nn.Hardtanh(0.0, 1.0))   # 3x64x64

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)          # hard-wired like vampvae.py:101
        return self.final_layer(self.decoder(h))

    def reparameterize(self, mu: Tensor, logvar: Tensor) -> Tensor:
        z, _, _ = F.reparam_kl(mu, logvar, want_kl=False)      # the closed-form KL is not part of this loss
        return z

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var, z]

    def prior_components(self, input: Tensor) -> List[Tensor]:
        """vampvae.py:144-149: encode the K embedded pseudo-inputs -> (prior_mu, prior_log_var), each [K, D]."""
        _, C, H, W = input.size()
        self.pseudo_input = self.pseudo_input.to(input.device)
        x = self.embed_pseudo(self.pseudo_input).view(-1, C, H, W)
        return self.encode(x)

    def loss_function(self, *args, **kwargs) -> dict:
        """vampvae.py:124-166: mse + M_N * (E_q log q - E_q log p_vamp)."""
        recons, input, mu, log_var, z = args[0], args[1], args[2], args[3], args[4]
        prior_mu, prior_log_var = self.prior_components(input)
        kl_rows = F.gauss_logq_rows(z, mu, log_var) - F.vamp_prior_rows(z, prior_mu, prior_log_var)
        loss, recon, _, neg_kld = F.elbo(recons, input, kl_rows, combine_mode=F.COMBINE_LINEAR, w=kwargs['M_N'])
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': neg_kld}
