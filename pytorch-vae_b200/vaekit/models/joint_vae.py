"""Drop-in for models/joint_vae.py: a Gaussian latent (K1) concatenated with one Gumbel-softmax
categorical latent (K8); the loss pushes both KL terms towards growing capacities (joint_vae.py:184-217)."""
from typing import List

import numpy as np
import torch
from torch import Tensor, nn

from .. import functional as F
from .base import GaussianConvVAE


class JointVAE(GaussianConvVAE):
    num_iter = 1

    def __init__(self, in_channels: int, latent_dim: int, categorical_dim: int, latent_min_capacity: float = 0.,
                 latent_max_capacity: float = 25., latent_gamma: float = 30., latent_num_iter: int = 25000,
                 categorical_min_capacity: float = 0., categorical_max_capacity: float = 25.,
                 categorical_gamma: float = 30., categorical_num_iter: int = 25000, hidden_dims: List = None,
                 temperature: float = 0.5, anneal_rate: float = 3e-5, anneal_interval: int = 100, alpha: float = 30.,
                 **kwargs) -> None:
        if hidden_dims is None:
            hidden_dims = [32, 64, 128, 256, 512]
        width = hidden_dims[-1]
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.categorical_dim = categorical_dim
        self.temp = temperature
        self.min_temp = temperature
        self.anneal_rate = anneal_rate
        self.anneal_interval = anneal_interval
        self.alpha = alpha
        self.cont_min, self.cont_max = latent_min_capacity, latent_max_capacity
        self.disc_min, self.disc_max = categorical_min_capacity, categorical_max_capacity
        self.cont_gamma, self.disc_gamma = latent_gamma, categorical_gamma
        self.cont_iter, self.disc_iter = latent_num_iter, categorical_num_iter
        self.fc_z = nn.Linear(width * 4, categorical_dim)                              # joint_vae.py:67
        self.decoder_input = nn.Linear(latent_dim + categorical_dim, width * 4)        # joint_vae.py:71
        self._module_order('encoder', 'fc_mu', 'fc_var', 'fc_z', 'decoder_input', 'decoder', 'final_layer')
        self.sampling_dist = torch.distributions.OneHotCategorical(1. / categorical_dim * torch.ones((categorical_dim, 1)))

    def encode(self, input: Tensor) -> List[Tensor]:
        h = torch.flatten(self.encoder(input), start_dim=1)
        return [self.fc_mu(h), self.fc_var(h), self.fc_z(h).view(-1, self.categorical_dim)]

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)          # hard-wired like joint_vae.py:137
        return self.final_layer(self.decoder(h))

    def reparameterize(self, mu: Tensor, log_var: Tensor, q: Tensor, eps: float = 1e-7) -> Tensor:
        """joint_vae.py:142-167: Gaussian sample (K1, KL rows kept for the loss) | Gumbel-softmax sample (K8)."""
        z = super().reparameterize(mu, log_var)
        s, _ = F.gumbel_softmax(q, self.temp, eps)
        return torch.cat([z, s.view(-1, self.categorical_dim)], dim=1)

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var, q = self.encode(input)
        z = self.reparameterize(mu, log_var, q)
        return [self.decode(z), input, q, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """joint_vae.py:174-222; num_iter advances in training mode only, the temperature anneals like cat_vae.py."""
        recons, input, q, mu, log_var = args[0], args[1], args[2], args[3], args[4]
        kld_weight = kwargs['M_N']
        batch_idx = kwargs['batch_idx']
        if batch_idx % self.anneal_interval == 0 and self.training:
            self.temp = np.maximum(self.temp * np.exp(- self.anneal_rate * batch_idx), self.min_temp)
        _, recons_loss, _, _ = F.elbo(recons, input)
        disc_curr = (self.disc_max - self.disc_min) * self.num_iter / float(self.disc_iter) + self.disc_min
        disc_curr = float(min(disc_curr, np.log(self.categorical_dim)))
        kld_disc_loss = torch.mean(F.cat_kl_rows(q.view(-1, 1, self.categorical_dim), 1e-7), dim=0)
        cont_curr = (self.cont_max - self.cont_min) * self.num_iter / float(self.cont_iter) + self.cont_min
        cont_curr = min(cont_curr, self.cont_max)
        kld_cont_loss = torch.mean(self._kl_rows(mu, log_var), dim=0)
        capacity_loss = self.disc_gamma * torch.abs(disc_curr - kld_disc_loss) + \
            self.cont_gamma * torch.abs(cont_curr - kld_cont_loss)
        loss = self.alpha * recons_loss + kld_weight * capacity_loss
        if self.training:
            self.num_iter += 1
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'Capacity_Loss': capacity_loss}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        z = torch.randn(num_samples, self.latent_dim)
        np_y = np.zeros((num_samples, self.categorical_dim), dtype=np.float32)
        np_y[range(num_samples), np.random.choice(self.categorical_dim, num_samples)] = 1
        z = torch.cat([z, torch.from_numpy(np_y)], dim=1).to(current_device)
        return self.decode(z)
