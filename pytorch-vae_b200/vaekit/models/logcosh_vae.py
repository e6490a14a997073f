"""Drop-in for models/logcosh_vae.py."""
from typing import List

from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE


class LogCoshVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, alpha: float = 100.,
                 beta: float = 10., **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.alpha = alpha
        self.beta = beta

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """logcosh_vae.py:125-155; the literal alpha*t + log(1+exp(-2 alpha t)) - log 2 form is kept,
        including its overflow to inf."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        kld_weight = kwargs['M_N']
        loss, recon, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var), recon_mode=F.RECON_LOGCOSH,
                                         alpha=self.alpha, combine_mode=F.COMBINE_LINEAR,
                                         w=self.beta * kld_weight)
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': neg_kld}
