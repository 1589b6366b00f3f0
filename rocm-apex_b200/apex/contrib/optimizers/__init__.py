from .distributed_fused_adam import DistributedFusedAdam  # noqa: F401
from .distributed_fused_lamb import DistributedFusedLAMB  # noqa: F401
