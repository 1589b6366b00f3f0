from .fused_adam import FusedAdam
from .fused_lamb import FusedLAMB
from .fused_sgd import FusedSGD
from .fused_mixed_precision_lamb import FusedMixedPrecisionLamb
from .fused_adagrad import FusedAdagrad
from .fused_novograd import FusedNovoGrad

__all__ = ["FusedAdam", "FusedLAMB", "FusedSGD", "FusedMixedPrecisionLamb", "FusedAdagrad", "FusedNovoGrad"]
