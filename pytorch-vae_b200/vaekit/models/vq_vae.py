"""Drop-in for models/vq_vae.py: VectorQuantizer, ResidualLayer, VQVAE (same constructor arguments,
parameter names, forward list and loss dict).  The conv stacks are ordinary torch modules."""
from typing import List, Union

from torch import Tensor, nn

from .. import functional as F
from .base import BaseVAE


class VectorQuantizer(nn.Module):
    """vq_vae.py:7-55.  forward() is ONE kernel (vk_vq_fwd): tcgen05 distance GEMM with the argmin,
    code gather, straight-through output and both mse terms fused in its epilogue; NCHW is read and
    written in place (no permute copies, no [R,K] distance or one-hot matrices)."""

    def __init__(self, num_embeddings: int, embedding_dim: int, beta: float = 0.25):
        super().__init__()
        self.K = num_embeddings
        self.D = embedding_dim
        self.beta = beta
        self.embedding = nn.Embedding(self.K, self.D)
        self.embedding.weight.data.uniform_(-1 / self.K, 1 / self.K)
        self.last_indices = None            # int32 [B*H*W]; not returned by the reference, kept for inspection

    def forward(self, latents: Tensor) -> Tensor:
        q, vq_loss, idx = F.vector_quantize(latents, self.embedding.weight, self.beta)
        self.last_indices = idx
        return q, vq_loss


class ResidualLayer(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.resblock = nn.Sequential(nn.Conv2d(in_channels, out_channels, kernel_size=3, padding=1, bias=False),
                                      nn.ReLU(True),
                                      nn.Conv2d(out_channels, out_channels, kernel_size=1, bias=False))

    def forward(self, input: Tensor) -> Tensor:
        return input + self.resblock(input)


def _conv_act(conv: nn.Module, act: nn.Module = None) -> nn.Sequential:
    return nn.Sequential(conv, act if act is not None else nn.LeakyReLU())


class VQVAE(BaseVAE):
    def __init__(self, in_channels: int, embedding_dim: int, num_embeddings: int, hidden_dims: List = None,
                 beta: float = 0.25, img_size: int = 64, **kwargs) -> None:
        super().__init__()
        self.embedding_dim = embedding_dim
        self.num_embeddings = num_embeddings
        self.img_size = img_size
        self.beta = beta
        if hidden_dims is None:
            hidden_dims = [128, 256]
        enc, c = [], in_channels
        for h in hidden_dims:
            enc.append(_conv_act(nn.Conv2d(c, out_channels=h, kernel_size=4, stride=2, padding=1)))
            c = h
        enc.append(_conv_act(nn.Conv2d(c, c, kernel_size=3, stride=1, padding=1)))
        enc += [ResidualLayer(c, c) for _ in range(6)]
        enc.append(nn.LeakyReLU())
        enc.append(_conv_act(nn.Conv2d(c, embedding_dim, kernel_size=1, stride=1)))
        self.encoder = nn.Sequential(*enc)
        self.vq_layer = VectorQuantizer(num_embeddings, embedding_dim, self.beta)
        top = hidden_dims[-1]
        dec = [_conv_act(nn.Conv2d(embedding_dim, top, kernel_size=3, stride=1, padding=1))]
        dec += [ResidualLayer(top, top) for _ in range(6)]
        dec.append(nn.LeakyReLU())
        hidden_dims.reverse()
        for a, b in zip(hidden_dims[:-1], hidden_dims[1:]):
            dec.append(_conv_act(nn.ConvTranspose2d(a, b, kernel_size=4, stride=2, padding=1)))
        dec.append(_conv_act(nn.ConvTranspose2d(hidden_dims[-1], out_channels=3, kernel_size=4, stride=2, padding=1),
                             nn.Tanh()))
        self.decoder = nn.Sequential(*dec)

    def encode(self, input: Tensor) -> List[Tensor]:
        return [self.encoder(input)]

    def decode(self, z: Tensor) -> Tensor:
        return self.decoder(z)

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        encoding = self.encode(input)[0]
        quantized_inputs, vq_loss = self.vq_layer(encoding)
        return [self.decode(quantized_inputs), input, vq_loss]

    def loss_function(self, *args, **kwargs) -> dict:
        """vq_vae.py:194-211: mse(recons, input) + vq_loss."""
        recons, input, vq_loss = args[0], args[1], args[2]
        _, recons_loss, _, _ = F.elbo(recons, input)
        loss = recons_loss + vq_loss
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'VQ_Loss': vq_loss}

    def sample(self, num_samples: int, current_device: Union[int, str], **kwargs) -> Tensor:
        raise Warning('VQVAE sampler is not implemented.')

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0]
