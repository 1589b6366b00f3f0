"""apex.parallel, restricted to the data-parallel gradient path (DistributedDataParallel, Reducer);
SyncBatchNorm and LARC are outside the multi-tensor optimizer path."""
from .distributed import DistributedDataParallel, Reducer, flat_dist_call  # noqa: F401
