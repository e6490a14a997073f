"""Drop-in for models/betatc_vae.py (Beta-TC-VAE, minibatch-stratified sampling)."""
import math
from typing import List

import torch
from torch import Tensor, nn

from .. import functional as F
from .base import BaseVAE, KLCache


class BetaTCVAE(BaseVAE):
    num_iter = 0        # class attribute shadowed per instance on the first training call (betatc_vae.py:10,197)

    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, anneal_steps: int = 200,
                 alpha: float = 1., beta: float = 6., gamma: float = 1., **kwargs) -> None:
        super().__init__()
        self.latent_dim = latent_dim
        self.anneal_steps = anneal_steps
        self.alpha = alpha
        self.beta = beta
        self.gamma = gamma
        if hidden_dims is None:
            hidden_dims = [32, 32, 32, 32]
        enc, c = [], in_channels
        for h in hidden_dims:
            enc.append(nn.Sequential(nn.Conv2d(c, out_channels=h, kernel_size=4, stride=2, padding=1), nn.LeakyReLU()))
            c = h
        self.encoder = nn.Sequential(*enc)
        self.fc = nn.Linear(hidden_dims[-1] * 16, 256)
        self.fc_mu = nn.Linear(256, latent_dim)
        self.fc_var = nn.Linear(256, latent_dim)
        self.decoder_input = nn.Linear(latent_dim, 256 * 2)
        hidden_dims.reverse()
        self.decoder = nn.Sequential(*[
            nn.Sequential(nn.ConvTranspose2d(a, b, kernel_size=3, stride=2, padding=1, output_padding=1), nn.LeakyReLU())
            for a, b in zip(hidden_dims[:-1], hidden_dims[1:])])
        last = hidden_dims[-1]
        self.final_layer = nn.Sequential(
            nn.ConvTranspose2d(last, last, kernel_size=3, stride=2, padding=1, output_padding=1), nn.LeakyReLU(),
            nn.Conv2d(last, out_channels=3, kernel_size=3, padding=1), nn.Tanh())
        self._kl = KLCache()

    def encode(self, input: Tensor) -> List[Tensor]:
        h = self.fc(torch.flatten(self.encoder(input), start_dim=1))
        return [self.fc_mu(h), self.fc_var(h)]

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 32, 4, 4)
        return self.final_layer(self.decoder(h))

    def reparameterize(self, mu: Tensor, logvar: Tensor) -> Tensor:
        """betatc_vae.py:115-125 as one kernel (vk_reparam_kl_fwd, KL side output not needed here)."""
        z, _, _ = F.reparam_kl(mu, logvar, want_kl=False)
        return z

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var, z]

    def log_density_gaussian(self, x: Tensor, mu: Tensor, logvar: Tensor):
        """betatc_vae.py:132-142.  Public helper of the reference; the loss does not go through it (the pairwise
        [B,B,D] density tensor is never materialised, see vk_tc_fwd) -- kept for callers, evaluated with torch ops."""
        norm = - 0.5 * (math.log(2 * math.pi) + logvar)
        return norm - 0.5 * ((x - mu) ** 2 * torch.exp(-logvar))

    def loss_function(self, *args, **kwargs) -> dict:
        """betatc_vae.py:144-211.  M_N is required (dataset_size = batch / M_N) although weight = 1."""
        recons, input, mu, log_var, z = args[0], args[1], args[2], args[3], args[4]
        M_N = kwargs['M_N']
        batch_size = z.shape[0]
        if self.training:
            self.num_iter += 1
            anneal_rate = min(0 + 1 * self.num_iter / self.anneal_steps, 1)
        else:
            anneal_rate = 1.
        _, recons_loss, _, _ = F.elbo(recons, input, recon_mode=F.RECON_MSE_SUM)
        latent, terms = F.betatc_terms(z, mu, log_var, M_N, self.alpha, self.beta, anneal_rate * self.gamma)
        loss = recons_loss / batch_size + latent
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'KLD': terms[0], 'TC_Loss': terms[1],
                'MI_Loss': terms[2]}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        z = torch.randn(num_samples, self.latent_dim)
        return self.decode(z.to(current_device))

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0]
