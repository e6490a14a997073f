"""Drop-in for models/mssim_vae.py: Gaussian VAE whose reconstruction term is 1 - MS-SSIM.
reparameterize + KL rows are K1; the multi-scale SSIM term is five depthwise-conv pyramid levels and stays torch
(cuDNN), as does the two-term combine.  Reference quirks kept on purpose so losses match bit for bit in fp32:
  * the window weights are exp(+(x - w//2)^2 / (2 sigma^2)), normalised (mssim_vae.py:206-209: no minus sign);
  * MSSIMVAE passes `size_average` into MSSIM's `normalize` slot (mssim_vae.py:80-82 vs :184-188);
  * the final product is prod(mcs[:-1]^w[:-1] * mssim[-1]^w[-1]) (mssim_vae.py:282)."""
from math import exp
from typing import List

import torch
from torch import Tensor, nn
from torch.nn import functional as TF

from .base import GaussianConvVAE

_LEVEL_WEIGHTS = (0.0448, 0.2856, 0.3001, 0.2363, 0.1333)


class MSSIM(nn.Module):
    def __init__(self, in_channels: int = 3, window_size: int = 11, normalize: bool = True,
                 size_average: bool = True) -> None:
        super().__init__()
        self.in_channels = in_channels
        self.window_size = window_size
        self.size_average = size_average
        self.normalize = normalize

    def gaussian_window(self, window_size: int, sigma: float) -> Tensor:
        half = window_size // 2
        taps = torch.tensor([exp((x - half) ** 2 / (2 * sigma ** 2)) for x in range(window_size)])
        return taps / taps.sum()

    def create_window(self, window_size: int, in_channels: int) -> Tensor:
        col = self.gaussian_window(window_size, 1.5).unsqueeze(1)
        plane = col.mm(col.t()).float()[None, None]
        return plane.expand(in_channels, 1, window_size, window_size).contiguous()

    def ssim(self, img1: Tensor, img2: Tensor, window_size: int, in_channel: int, size_average: bool):
        window = self.create_window(window_size, in_channel).to(img1.device)

        def blur(t):
            return TF.conv2d(t, window, padding=window_size // 2, groups=in_channel)

        mu1, mu2 = blur(img1), blur(img2)
        mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
        sigma1_sq = blur(img1 * img1) - mu1_sq
        sigma2_sq = blur(img2 * img2) - mu2_sq
        sigma12 = blur(img1 * img2) - mu1_mu2
        C1, C2 = 0.01 ** 2, 0.03 ** 2                       # dynamic range fixed to 1, mssim_vae.py:232-234
        v1 = 2.0 * sigma12 + C2
        v2 = sigma1_sq + sigma2_sq + C2
        cs = torch.mean(v1 / v2)
        ssim_map = ((2 * mu1_mu2 + C1) * v1) / ((mu1_sq + mu2_sq + C1) * v2)
        ret = ssim_map.mean() if size_average else ssim_map.mean(1).mean(1).mean(1)
        return ret, cs

    def forward(self, img1: Tensor, img2: Tensor) -> Tensor:
        weights = torch.tensor(_LEVEL_WEIGHTS, dtype=torch.float32).to(img1.device)
        mssim, mcs = [], []
        for _ in range(len(_LEVEL_WEIGHTS)):
            sim, cs = self.ssim(img1, img2, self.window_size, self.in_channels, self.size_average)
            mssim.append(sim)
            mcs.append(cs)
            img1 = TF.avg_pool2d(img1, (2, 2))
            img2 = TF.avg_pool2d(img2, (2, 2))
        mssim, mcs = torch.stack(mssim), torch.stack(mcs)
        if self.normalize:
            mssim, mcs = (mssim + 1) / 2, (mcs + 1) / 2
        pow1, pow2 = mcs ** weights, mssim ** weights
        return 1 - torch.prod(pow1[:-1] * pow2[-1])


class MSSIMVAE(GaussianConvVAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, window_size: int = 11,
                 size_average: bool = True, **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.in_channels = in_channels
        self.mssim_loss = MSSIM(in_channels, window_size, size_average)     # positional, like mssim_vae.py:80-82

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)          # hard-wired like mssim_vae.py:108
        return self.final_layer(self.decoder(h))

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        mu, log_var = self.encode(input)
        z = self.reparameterize(mu, log_var)
        return [self.decode(z), input, mu, log_var]

    def loss_function(self, *args, **kwargs) -> dict:
        """mssim_vae.py:131-153: (1 - MS-SSIM) + M_N * mean KL; 'KLD' is the negated mean and keeps its graph."""
        recons, input, mu, log_var = args[0], args[1], args[2], args[3]
        recons_loss = self.mssim_loss(recons, input)
        kld_loss = self._kl_rows(mu, log_var).mean(dim=0)
        return {'loss': recons_loss + kwargs['M_N'] * kld_loss, 'Reconstruction_Loss': recons_loss, 'KLD': -kld_loss}
