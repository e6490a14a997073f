"""Abstract model interface, mirroring models/base.py:5-28 of the reference, plus
the conv encoder/decoder builders shared by the drop-in classes.  The conv
stacks are ordinary torch modules (out of scope for the kernels) with the
reference's parameter names, so its checkpoints load unchanged."""
from abc import abstractmethod
from typing import Any, List

import torch
from torch import Tensor, nn


class BaseVAE(nn.Module):
    def __init__(self) -> None:
        super().__init__()

    def encode(self, input: Tensor) -> List[Tensor]:
        raise NotImplementedError

    def decode(self, input: Tensor) -> Any:
        raise NotImplementedError

    def sample(self, batch_size: int, current_device: int, **kwargs) -> Tensor:
        raise NotImplementedError

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        raise NotImplementedError

    def _module_order(self, *names: str) -> None:
        """Registration order of the sub-modules = the reference's: `state_dict()` and `parameters()` then enumerate in
        the reference's order, which is what an optimizer state saved by run.py:80-99 is indexed by."""
        mods = self._modules
        self._modules = type(mods)([(n, mods[n]) for n in names] + [(n, m) for n, m in mods.items() if n not in names])

    @abstractmethod
    def forward(self, *inputs: Tensor) -> Tensor:
        pass

    @abstractmethod
    def loss_function(self, *inputs: Any, **kwargs) -> Tensor:
        pass


def _block(conv: nn.Module, channels: int, batch_norm: bool) -> nn.Sequential:
    layers = [conv] + ([nn.BatchNorm2d(channels)] if batch_norm else []) + [nn.LeakyReLU()]
    return nn.Sequential(*layers)


def down_stack(in_channels: int, widths, kernel=3, stride=2, padding=1, batch_norm=True) -> nn.Sequential:
    blocks, c = [], in_channels
    for w in widths:
        blocks.append(_block(nn.Conv2d(c, w, kernel_size=kernel, stride=stride, padding=padding), w, batch_norm))
        c = w
    return nn.Sequential(*blocks)


def up_stack(widths, batch_norm=True) -> nn.Sequential:
    """widths already in decoder order (widest first)."""
    return nn.Sequential(*[
        _block(nn.ConvTranspose2d(a, b, kernel_size=3, stride=2, padding=1, output_padding=1), b, batch_norm)
        for a, b in zip(widths[:-1], widths[1:])])


def output_head(channels: int, batch_norm=True, out_channels=3) -> nn.Sequential:
    layers = [nn.ConvTranspose2d(channels, channels, kernel_size=3, stride=2, padding=1, output_padding=1)]
    if batch_norm:
        layers.append(nn.BatchNorm2d(channels))
    layers += [nn.LeakyReLU(), nn.Conv2d(channels, out_channels, kernel_size=3, padding=1), nn.Tanh()]
    return nn.Sequential(*layers)


class KLCache:
    """Hands the KL rows computed inside the fused reparameterize kernel to the
    loss_function that is later called with the *same* mu / log_var tensors
    (identity and in-place version checked); anything else recomputes them."""

    def __init__(self):
        self._entry = None

    def put(self, mu, log_var, kl):
        self._entry = (mu, log_var, mu._version, log_var._version, kl)

    def take(self, mu, log_var):
        e, self._entry = self._entry, None
        if e is not None and e[0] is mu and e[1] is log_var and e[2] == mu._version and e[3] == log_var._version:
            return e[4]
        return None


class GaussianConvVAE(BaseVAE):
    """Encoder -> (mu, log_var) -> decoder skeleton of vanilla_vae.py:10-105 (identical in
    beta_vae / logcosh_vae / iwae / miwae): 64x64 inputs, five stride-2 convs."""

    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None) -> None:
        super().__init__()
        self.latent_dim = latent_dim
        if hidden_dims is None:
            hidden_dims = [32, 64, 128, 256, 512]
        self.encoder = down_stack(in_channels, hidden_dims)
        self.fc_mu = nn.Linear(hidden_dims[-1] * 4, latent_dim)
        self.fc_var = nn.Linear(hidden_dims[-1] * 4, latent_dim)
        self.decoder_input = nn.Linear(latent_dim, hidden_dims[-1] * 4)
        hidden_dims.reverse()                      # the reference reverses the caller's list in place too
        self._top = hidden_dims[0]
        self.decoder = up_stack(hidden_dims)
        self.final_layer = output_head(hidden_dims[-1])
        self._kl = KLCache()

    def encode(self, input: Tensor) -> List[Tensor]:
        h = torch.flatten(self.encoder(input), start_dim=1)
        return [self.fc_mu(h), self.fc_var(h)]

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, self._top, 2, 2)
        return self.final_layer(self.decoder(h))

    def reparameterize(self, mu: Tensor, logvar: Tensor) -> Tensor:
        """vanilla_vae.py:107-117 as ONE kernel (vk_reparam_kl_fwd): in-register Philox noise,
        z written once, KL rows emitted on the side for the coming loss_function call."""
        from .. import functional as F
        z, kl, _ = F.reparam_kl(mu, logvar)
        self._kl.put(mu, logvar, kl)
        return z

    def _kl_rows(self, mu: Tensor, log_var: Tensor) -> Tensor:
        from .. import functional as F
        kl = self._kl.take(mu, log_var)
        return kl if kl is not None else F.kl_rows(mu, log_var)

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        z = torch.randn(num_samples, self.latent_dim)       # CPU generator, as vanilla_vae.py:158-161
        return self.decode(z.to(current_device))

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0]


class DeterministicConvAE(BaseVAE):
    """Encoder -> z -> decoder skeleton shared by wae_mmd.py:10-105 and swae.py:10-108 (fc_z instead of
    fc_mu / fc_var; the decoder reshape is hard-wired to 512 channels as in the reference)."""

    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None) -> None:
        super().__init__()
        self.latent_dim = latent_dim
        if hidden_dims is None:
            hidden_dims = [32, 64, 128, 256, 512]
        self.encoder = down_stack(in_channels, hidden_dims)
        self.fc_z = nn.Linear(hidden_dims[-1] * 4, latent_dim)
        self.decoder_input = nn.Linear(latent_dim, hidden_dims[-1] * 4)
        hidden_dims.reverse()
        self.decoder = up_stack(hidden_dims)
        self.final_layer = output_head(hidden_dims[-1])

    def encode(self, input: Tensor) -> Tensor:
        return self.fc_z(torch.flatten(self.encoder(input), start_dim=1))

    def decode(self, z: Tensor) -> Tensor:
        h = self.decoder_input(z).view(-1, 512, 2, 2)
        return self.final_layer(self.decoder(h))

    def forward(self, input: Tensor, **kwargs) -> List[Tensor]:
        z = self.encode(input)
        return [self.decode(z), input, z]

    def sample(self, num_samples: int, current_device: int, **kwargs) -> Tensor:
        z = torch.randn(num_samples, self.latent_dim)
        return self.decode(z.to(current_device))

    def generate(self, x: Tensor, **kwargs) -> Tensor:
        return self.forward(x)[0]
