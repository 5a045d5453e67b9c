"""vgan_b200 — B200-native (sm_100a) squared-MMD forward+backward behind V-GAN's loss interface.

Public surface mirrors the reference (/root/reference/src): ``RBF``, ``MMDLossConstrained``
(models/Mmd_loss_constrained.py), ``MMDStatistic`` (the torch_two_sample call V-GAN relies on),
``VGAN`` / ``VGAN_no_kl`` (vgan.py)."""
from .loss import RBF, MMDLossConstrained, mmd2_eval  # noqa: F401
from .statistic import MMDStatistic  # noqa: F401
from .subspace import unique_subspaces  # noqa: F401
from .vgan import VGAN, VGAN_no_kl  # noqa: F401
from . import ops  # noqa: F401  (registers torch.ops.vgan_b200.*)

__all__ = ["RBF", "MMDLossConstrained", "MMDStatistic", "VGAN", "VGAN_no_kl", "mmd2_eval", "unique_subspaces"]
