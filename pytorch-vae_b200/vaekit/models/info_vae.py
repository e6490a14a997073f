"""Drop-in for models/info_vae.py: a Gaussian encoder whose loss adds an un-weighted MMD
between the sampled z and a prior draw to the scaled ELBO terms.  No new kernels: K1 (reparameterize + KL rows),
K2 (scaled combine) and K4 (the WAE Gram sweep)."""
from typing import List

from torch import Tensor

from .. import functional as F
from .base import GaussianConvVAE
from .wae_mmd import KernelMatrixMixin


class InfoVAE(KernelMatrixMixin, GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, alpha: float = -0.5,
                 beta: float = 5.0, reg_weight: int = 100, kernel_type: str = 'imq', latent_var: float = 2.,
                 **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.reg_weight = reg_weight
        self.kernel_type = kernel_type
        self.z_var = latent_var
        assert alpha <= 0, 'alpha must be negative or zero.'
        self.alpha = alpha
        self.beta = beta
        self.mmd_passes = 0          # 0: auto (sample-checked reduced passes where safe), 3: parity sweep, 2 / 1: reduced

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)          # hard-wired like info_vae.py:106
        return self.final_layer(self.decoder(h))

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, z, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """info_vae.py:117-134: beta * mse + (1 - alpha) * M_N * kld + (alpha + reg_weight - 1) / (B (B-1)) * mmd.
        'Reconstruction_Loss', 'MMD' and 'KLD' (= -kld) keep their graph like the reference's."""
        recons, input, z, mu, log_var = args[0], args[1], args[2], args[3], args[4]
        batch_size = input.size(0)
        bias_corr = batch_size * (batch_size - 1)
        kld_weight = kwargs['M_N']
        elbo, recons_loss, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var),
                                               combine_mode=F.COMBINE_SCALED, C=self.beta,
                                               w=(1. - self.alpha) * kld_weight)
        mmd_loss = self.compute_mmd(z)
        loss = elbo + (self.alpha + self.reg_weight - 1.) / bias_corr * mmd_loss
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'MMD': mmd_loss, 'KLD': neg_kld}

    def compute_mmd(self, z: Tensor) -> Tensor:
        """info_vae.py:203-215: prior_z from the contract Philox stream, kernel means / off-diagonal sums with
        weight 1 (the scaling happens in loss_function)."""
        if self.kernel_type not in F.KERNEL_TYPES:
            raise ValueError('Undefined kernel type.')
        prior_z = F.philox_normal(z.shape, z.device)
        self.last_prior_z = prior_z
        mmd, terms = F.mmd(z, prior_z, 1.0, self.kernel_type, self.z_var, passes=self.mmd_passes)
        self.last_terms = terms
        return mmd
