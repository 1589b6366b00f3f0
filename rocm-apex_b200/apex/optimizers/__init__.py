from .fused_adam import FusedAdam
from .fused_lamb import FusedLAMB
from .fused_sgd import FusedSGD
from .fused_mixed_precision_lamb import FusedMixedPrecisionLamb

__all__ = ["FusedAdam", "FusedLAMB", "FusedSGD", "FusedMixedPrecisionLamb"]
