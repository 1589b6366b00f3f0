"""apex.amp, restricted to the loss-scaling runtime that drives the fused kernels."""
from ._amp_state import master_params, _amp_state  # noqa: F401
from .frontend import initialize, state_dict, load_state_dict  # noqa: F401
from .handle import scale_loss  # noqa: F401
from .scaler import LossScaler  # noqa: F401
