"""Drop-in for models/cat_vae.py: categorical VAE with the Gumbel-softmax reparameterisation.
reparameterize is one kernel (vk_gumbel_softmax_fwd: Philox uniforms + Gumbel noise + tempered softmax), the KL against
the uniform categorical prior is one kernel (vk_cat_kl_rows_fwd), the reconstruction term and the combine are K2."""
from typing import List

import numpy as np
import torch
from torch import Tensor, nn

from .. import functional as F
from .base import BaseVAE, down_stack, output_head, up_stack


class CategoricalVAE(BaseVAE):
    def __init__(self, in_channels: int, latent_dim: int, categorical_dim: int = 40, hidden_dims: List = None,
                 temperature: float = 0.5, anneal_rate: float = 3e-5, anneal_interval: int = 100, alpha: float = 30.,
                 **kwargs) -> None:
        super().__init__()
        self.latent_dim = latent_dim
        self.categorical_dim = categorical_dim
        self.temp = temperature
        self.min_temp = temperature
        self.anneal_rate = anneal_rate
        self.anneal_interval = anneal_interval
        self.alpha = alpha
        if hidden_dims is None:
            hidden_dims = [32, 64, 128, 256, 512]
        self.encoder = down_stack(in_channels, hidden_dims)
        self.fc_z = nn.Linear(hidden_dims[-1] * 4, latent_dim * categorical_dim)
        self.decoder_input = nn.Linear(latent_dim * categorical_dim, hidden_dims[-1] * 4)
        hidden_dims.reverse()
        self.decoder = up_stack(hidden_dims)
        self.final_layer = output_head(hidden_dims[-1])
        self.sampling_dist = torch.distributions.OneHotCategorical(1. / categorical_dim * torch.ones((categorical_dim, 1)))

    def encode(self, input: Tensor) -> List[Tensor]:
        h = torch.flatten(self.encoder(input), start_dim=1)
        return [self.fc_z(h).view(-1, self.latent_dim, self.categorical_dim)]

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)          # hard-wired like cat_vae.py:111
        return self.final_layer(self.decoder(h))

    def reparameterize(self, z: Tensor, eps: float = 1e-7) -> Tensor:
        """cat_vae.py:118-131: [B, D, Q] logits -> [B, D * Q] Gumbel-softmax sample, one kernel."""
        s, _ = F.gumbel_softmax(z, self.temp, eps)
        return s.view(-1, self.latent_dim * self.categorical_dim)

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        q = self.encode(input)[0]
        z = self.reparameterize(q)
        return [self.decode(z), input, q]

    def loss_function(self, *args, **kwargs) -> dict:
        """cat_vae.py:133-173: alpha * mse + M_N * KL(softmax(q) || uniform); the temperature is annealed every
        anneal_interval batches while training (host state, like the reference)."""
        recons, input, q = args[0], args[1], args[2]
        kld_weight = kwargs['M_N']
        batch_idx = kwargs['batch_idx']
        if batch_idx % self.anneal_interval == 0 and self.training:
            self.temp = np.maximum(self.temp * np.exp(- self.anneal_rate * batch_idx), self.min_temp)
        loss, recons_loss, _, neg_kld = F.elbo(recons, input, F.cat_kl_rows(q, 1e-7), combine_mode=F.COMBINE_SCALED,
                                               C=self.alpha, w=kld_weight)
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'KLD': neg_kld}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        M = num_samples * self.latent_dim                       # one-hot draws on the numpy generator, cat_vae.py:186-192
        np_y = np.zeros((M, self.categorical_dim), dtype=np.float32)
        np_y[range(M), np.random.choice(self.categorical_dim, M)] = 1
        np_y = np.reshape(np_y, [M // self.latent_dim, self.latent_dim, self.categorical_dim])
        z = torch.from_numpy(np_y).view(num_samples, self.latent_dim * self.categorical_dim).to(current_device)
        return self.decode(z)

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0]
