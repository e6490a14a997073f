"""Drop-in for models/dfcvae.py: deep-feature-consistent VAE.  reparameterize + KL rows are K1;
the pixel term and each of the four feature-map terms are K2 reductions (one launch per pair instead of torch's
sub / pow / mean chain); the frozen VGG19-BN feature extractor stays a torch module.

The reference downloads ImageNet weights in its constructor (dfcvae.py:81).  There is no network on the target boxes,
so the extractor can be handed in (`feature_network=`: any module with a `.features` container whose children are
addressed by the reference's layer ids) or built without weights (`pretrained=False`); the default keeps the
reference's behaviour and fails with a plain message when the download is impossible."""
from typing import List

import torch
from torch import Tensor, nn

from .. import functional as F
from .base import GaussianConvVAE


def _vgg19_bn(pretrained: bool) -> nn.Module:
    from torchvision.models import vgg19_bn
    if not pretrained:
        return vgg19_bn(weights=None)
    try:
        return vgg19_bn(weights="IMAGENET1K_V1")
    except Exception as e:                                       # no egress / no cached checkpoint
        raise RuntimeError("vaekit: DFCVAE needs the ImageNet vgg19_bn weights (dfcvae.py:81) and they could not be "
                           "loaded (%s); pass feature_network=<module> or pretrained=False" % (e,)) from e


class DFCVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, alpha: float = 1, beta: float = 0.5,
                 feature_network: nn.Module = None, pretrained: bool = True, **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.alpha = alpha
        self.beta = beta
        self.feature_network = feature_network if feature_network is not None else _vgg19_bn(pretrained)
        for param in self.feature_network.parameters():          # frozen, dfcvae.py:84-87
            param.requires_grad = False
        self.feature_network.eval()

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)            # hard-wired like dfcvae.py:118
        return self.final_layer(self.decoder(h))

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        recons = self.decode(z)
        return [recons, input, self.extract_features(recons), self.extract_features(input), mu, log_var]

    def extract_features(self, input: Tensor, feature_layers: List = None) -> List[Tensor]:
        """dfcvae.py:142-162: run the extractor's `.features` children in order, keep the outputs of the named ones."""
        if feature_layers is None:
            feature_layers = ['14', '24', '34', '43']
        features, h = [], input
        for key, module in self.feature_network.features._modules.items():
            h = module(h)
            if key in feature_layers:
                features.append(h)
        return features

    def loss_function(self, *args, **kwargs) -> dict:
        """dfcvae.py:164-190: beta * (mse + sum of feature mse) + alpha * M_N * mean KL."""
        recons, input, recons_features, input_features, mu, log_var = args[:6]
        _, recons_loss, _, neg_kld = F.elbo(recons, input, self._kl_rows(mu, log_var))
        feature_loss = 0.0
        for r, i in zip(recons_features, input_features):
            feature_loss = feature_loss + F.elbo(r, i)[1]
        loss = self.beta * (recons_loss + feature_loss) - self.alpha * kwargs['M_N'] * neg_kld
        return {'loss': loss, 'Reconstruction_Loss': self.beta * recons_loss, 'KLD': neg_kld,
                'feature_loss': self.beta * feature_loss}

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        z = torch.randn(num_samples, self.latent_dim)
        return self.decode(z.to(current_device))
