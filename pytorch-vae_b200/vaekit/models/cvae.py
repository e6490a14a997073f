"""Drop-in for models/cvae.py: class-conditional Gaussian VAE.  The label embedding, the 1x1 data
embedding and the conv stacks stay torch modules; reparameterize + KL rows are K1, the loss is K2."""
from typing import List

import torch
from torch import Tensor, nn

from .. import functional as F
from .base import GaussianConvVAE, down_stack


class ConditionalVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, num_classes: int, latent_dim: int, hidden_dims: List = None,
                 img_size: int = 64, **kwargs) -> None:
        if hidden_dims is None:
            hidden_dims = [32, 64, 128, 256, 512]
        widths = list(hidden_dims)
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.img_size = img_size
        self.embed_class = nn.Linear(num_classes, img_size * img_size)            # cvae.py:23
        self.embed_data = nn.Conv2d(in_channels, in_channels, kernel_size=1)      # cvae.py:24
        self.encoder = down_stack(in_channels + 1, widths)                        # extra label channel, cvae.py:30
        self.decoder_input = nn.Linear(latent_dim + num_classes, widths[-1] * 4)  # cvae.py:51
        self._module_order('embed_class', 'embed_data', 'encoder', 'fc_mu', 'fc_var', 'decoder_input', 'decoder',
                           'final_layer')

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)          # hard-wired like cvae.py:102
        return self.final_layer(self.decoder(h))

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        y = kwargs['labels'].float()
        embedded_class = self.embed_class(y).view(-1, self.img_size, self.img_size).unsqueeze(1)
        x = torch.cat([self.embed_data(input), embedded_class], dim=1)
        mu, log_var = self.encode(x)
        z = self.reparameterize(mu, log_var)
        return [self.decode(torch.cat([z, y], dim=1)), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """cvae.py:131-146: mse + M_N * KL; unlike VanillaVAE neither side output is detached."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        loss, recon, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var),
                                         combine_mode=F.COMBINE_LINEAR, w=kwargs['M_N'])
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': neg_kld}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        y = kwargs['labels'].float()
        z = torch.randn(num_samples, self.latent_dim).to(current_device)          # CPU generator, cvae.py:161-164
        return self.decode(torch.cat([z, y], dim=1))

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x, **kwargs)[0]
