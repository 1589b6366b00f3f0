from .distributed_fused_adam import DistributedFusedAdam  # noqa: F401
