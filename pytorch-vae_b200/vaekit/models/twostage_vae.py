"""Drop-in for models/twostage_vae.py.  The first-stage loss is the vanilla ELBO with neither side
output detached (twostage_vae.py:141-167) on K1 / K2; the second-stage MLP encoder / decoder are built with the
reference's module names so checkpoints load, and like in the reference nothing calls them."""
from typing import List

from torch import Tensor, nn

from .. import functional as F
from .base import GaussianConvVAE


def _mlp(in_features: int, widths) -> nn.Sequential:
    blocks = []
    for w in widths:
        blocks.append(nn.Sequential(nn.Linear(in_features, w), nn.BatchNorm1d(w), nn.LeakyReLU()))
        in_features = w
    return nn.Sequential(*blocks)


class TwoStageVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, hidden_dims2: List = None,
                 **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        if hidden_dims2 is None:
            hidden_dims2 = [1024, 1024]
        self.encoder2 = _mlp(latent_dim, hidden_dims2)                      # twostage_vae.py:76-86
        self.fc_mu2 = nn.Linear(hidden_dims2[-1], latent_dim)
        self.fc_var2 = nn.Linear(hidden_dims2[-1], latent_dim)
        hidden_dims2.reverse()                                              # in place, like the reference
        self.decoder2 = _mlp(latent_dim, hidden_dims2)

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        loss, recon, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var),
                                         combine_mode=F.COMBINE_LINEAR, w=kwargs['M_N'])
        return {'loss': loss, 'Reconstruction_Loss': recon, 'KLD': neg_kld}
