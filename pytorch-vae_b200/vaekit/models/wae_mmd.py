"""Drop-in for models/wae_mmd.py (WAE with an MMD penalty, RBF or IMQ kernel)."""
from typing import List

from torch import Tensor

from .. import functional as F
from .base import DeterministicConvAE


class KernelMatrixMixin:
    """compute_kernel / compute_rbf / compute_inv_mult_quad with the reference's signatures and results
    (wae_mmd.py:125-191, info_vae.py:150-201) on vk_pair_kernel_*.  compute_mmd does not go through them: it never
    needs the [N, N] matrices (vk_mmd_fwd)."""

    def compute_kernel(self, x1: Tensor, x2: Tensor) -> Tensor:
        D = x1.size(1)
        N = x1.size(0)
        x1 = x1.unsqueeze(-2).expand(N, N, D)      # column tensor, wae_mmd.py:132,140
        x2 = x2.unsqueeze(-3).expand(N, N, D)      # row tensor, wae_mmd.py:133,141
        if self.kernel_type == 'rbf':
            return self.compute_rbf(x1, x2)
        elif self.kernel_type == 'imq':
            return self.compute_inv_mult_quad(x1, x2)
        raise ValueError('Undefined kernel type.')

    def compute_rbf(self, x1: Tensor, x2: Tensor, eps: float = 1e-7) -> Tensor:
        return F.pair_kernel(x1, x2, 'rbf', self.z_var, eps)

    def compute_inv_mult_quad(self, x1: Tensor, x2: Tensor, eps: float = 1e-7) -> Tensor:
        return F.pair_kernel(x1, x2, 'imq', self.z_var, eps)


class WAE_MMD(KernelMatrixMixin, DeterministicConvAE):
    def __init__(self, in_channels: int, latent_dim: int, hidden_dims: List = None, reg_weight: int = 100,
                 kernel_type: str = 'imq', latent_var: float = 2., **kwargs) -> None:
        super().__init__(in_channels, latent_dim, hidden_dims)
        self.reg_weight = reg_weight
        self.kernel_type = kernel_type
        self.z_var = latent_var
        self.mmd_passes = 0          # 0: auto (sample-checked reduced passes where safe), 3: parity sweep, 2 / 1: reduced

    def loss_function(self, *args, **kwargs) -> dict:
        """wae_mmd.py:107-123 (does not read M_N; raises ZeroDivisionError at batch size 1 like the reference)."""
        recons, input, z = args[0], args[1], args[2]
        batch_size = input.size(0)
        bias_corr = batch_size * (batch_size - 1)
        reg_weight = self.reg_weight / bias_corr
        _, recons_loss, _, _ = F.elbo(recons, input)
        mmd_loss = self.compute_mmd(z, reg_weight)
        loss = recons_loss + mmd_loss
        return {'loss': loss, 'Reconstruction_Loss': recons_loss, 'MMD': mmd_loss}

    def compute_mmd(self, z: Tensor, reg_weight: float) -> Tensor:
        """wae_mmd.py:193-204: prior_z is drawn from the contract Philox stream (torch.manual_seed governs it);
        the three kernel matrices are never materialised (vk_mmd_fwd)."""
        if self.kernel_type not in F.KERNEL_TYPES:
            raise ValueError('Undefined kernel type.')
        prior_z = F.philox_normal(z.shape, z.device)
        self.last_prior_z = prior_z
        mmd, terms = F.mmd(z, prior_z, reg_weight, self.kernel_type, self.z_var, passes=self.mmd_passes)
        self.last_terms = terms       # {pp, zz, pz} kernel means (rbf) / off-diagonal sums (imq), for inspection
        return mmd
