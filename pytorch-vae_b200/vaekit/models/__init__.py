"""Model registry with the reference's keys (models/__init__.py:65-91) for the classes on the
hot path, plus the SURVEY 8(f) "next" rows built so far: InfoVAE (rank 1) and the Gaussian-KL models ConditionalVAE, DIPVAE,
FactorVAE, TwoStageVAE, MSSIMVAE and DFCVAE (rank 2), CategoricalVAE and JointVAE (rank 4, Gumbel-softmax kernels) and
VampVAE (rank 4, VampPrior kernels); the remaining reference variants (GammaVAE, HVAE, LVAE) are out of scope."""
from .base import BaseVAE
from .vanilla_vae import VanillaVAE
from .beta_vae import BetaVAE
from .logcosh_vae import LogCoshVAE
from .iwae import IWAE
from .miwae import MIWAE
from .vq_vae import VQVAE, VectorQuantizer
from .wae_mmd import WAE_MMD
from .info_vae import InfoVAE
from .swae import SWAE
from .betatc_vae import BetaTCVAE
from .cvae import ConditionalVAE
from .dip_vae import DIPVAE
from .fvae import FactorVAE
from .cat_vae import CategoricalVAE
from .joint_vae import JointVAE
from .twostage_vae import TwoStageVAE
from .mssim_vae import MSSIMVAE
from .dfcvae import DFCVAE
from .vampvae import VampVAE

VAE = VanillaVAE
GaussianVAE = VanillaVAE
CVAE = ConditionalVAE
GumbelVAE = CategoricalVAE

vae_models = {
    'IWAE': IWAE,
    'MIWAE': MIWAE,
    'BetaVAE': BetaVAE,
    'LogCoshVAE': LogCoshVAE,
    'VanillaVAE': VanillaVAE,
    'VQVAE': VQVAE,
    'WAE_MMD': WAE_MMD,
    'InfoVAE': InfoVAE,
    'SWAE': SWAE,
    'BetaTCVAE': BetaTCVAE,
    'ConditionalVAE': ConditionalVAE,
    'DIPVAE': DIPVAE,
    'FactorVAE': FactorVAE,
    'CategoricalVAE': CategoricalVAE,
    'JointVAE': JointVAE,
    'TwoStageVAE': TwoStageVAE,
    'MSSIMVAE': MSSIMVAE,
    'DFCVAE': DFCVAE,
    'VampVAE': VampVAE,
}
