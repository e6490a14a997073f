"""Drop-in for models/dip_vae.py: DIP-VAE II.  Summed reconstruction and KL terms are K2 / K1; the
[D, D] covariance regulariser of dip_vae.py:147-158 is a tiny torch computation on mu / log_var and stays torch."""
from typing import List

import torch
from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE


class DIPVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, lambda_diag: float = 10.,
                 lambda_offdiag: float = 5., **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.lambda_diag = lambda_diag
        self.lambda_offdiag = lambda_offdiag

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """dip_vae.py:124-164: mse SUM + M_N * (KL summed over the batch) + DIP loss."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        kld_weight = kwargs['M_N']
        B = mu.shape[0]
        # K2 combines with the batch MEAN of the KL rows: sum = B * mean
        elbo, recons_loss, _, neg_kld_mean = F.elbo(recons, input, self._kl_rows(mu, log_var), recon_mode=F.RECON_MSE_SUM,
                                                    combine_mode=F.COMBINE_LINEAR, w=kld_weight * B)
        centered_mu = mu - mu.mean(dim=1, keepdim=True)
        cov_mu = centered_mu.t().matmul(centered_mu).squeeze()
        cov_z = cov_mu + torch.mean(torch.diagonal((2. * log_var).exp(), dim1=0), dim=0)      # DIP-VAE II, dip_vae.py:151
        cov_diag = torch.diag(cov_z)
        cov_offdiag = cov_z - torch.diag(cov_diag)
        dip_loss = self.lambda_offdiag * torch.sum(cov_offdiag ** 2) + self.lambda_diag * torch.sum((cov_diag - 1) ** 2)
        return {'loss': elbo + dip_loss, 'Reconstruction_Loss': recons_loss, 'KLD': neg_kld_mean * B, 'DIP_Loss': dip_loss}
