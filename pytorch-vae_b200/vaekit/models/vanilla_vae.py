"""Drop-in for models/vanilla_vae.py: same constructor, forward list and loss dict."""
from typing import List

from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE


class VanillaVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        # extension (the reference has no BCE term anywhere; its constructor swallows unknown kwargs):
        # recons_loss='bce' makes loss_function use F.binary_cross_entropy(recons, input) -- recons must lie in [0, 1],
        # i.e. the caller's decoder must not end in the reference's Tanh
        self.recons_loss = kwargs.get('recons_loss', 'mse')
        if self.recons_loss not in ('mse', 'bce'):
            raise ValueError('Undefined reconstruction loss.')

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """vanilla_vae.py:124-146: mse + M_N * KL in one streaming kernel (vk_elbo_fwd)."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        kld_weight = kwargs['M_N']
        loss, recon, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var),
                                         recon_mode=F.RECON_BCE_MEAN if self.recons_loss == 'bce' else F.RECON_MSE_MEAN,
                                         combine_mode=F.COMBINE_LINEAR, w=kld_weight)
        return {'loss': loss, 'Reconstruction_Loss': recon.detach(), 'KLD': neg_kld.detach()}
