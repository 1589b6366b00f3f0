from .distributed_fused_adam import DistributedFusedAdam, reshard_state_dicts  # noqa: F401
from .distributed_fused_lamb import DistributedFusedLAMB  # noqa: F401
