"""Host-side mirror of the reference's Python for the multi-tensor fused-optimizer path only
(apex.multi_tensor_apply, apex.optimizers, apex.amp's loss-scaling runtime, and
apex.contrib.optimizers.DistributedFusedAdam).  Same names, constructor arguments, state keys and
error behaviour as the reference; everything else of apex is out of scope."""
from . import multi_tensor_apply  # noqa: F401
from . import optimizers  # noqa: F401
from . import amp  # noqa: F401
