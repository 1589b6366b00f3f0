"""Process-wide amp state (reference: apex/amp/_amp_state.py)."""
import torch


class AmpState(object):
    def __init__(self):
        self.hard_override = False
        self.allow_incoming_model_not_fp32 = False
        self.verbosity = 1
        self.loss_scalers = []
        self.min_loss_scale = None
        self.max_loss_scale = 2. ** 24


_amp_state = AmpState()


def warn_or_err(msg):
    if _amp_state.hard_override:
        print("Warning:  " + msg)
    else:
        raise RuntimeError(msg)


def maybe_print(msg, rank0=False):
    """Prints when verbosity > 0; with rank0=True only on rank 0 of an initialised process group."""
    if _amp_state.verbosity <= 0:
        return
    if rank0 and torch.distributed.is_available() and torch.distributed.is_initialized() \
            and torch.distributed.get_rank() != 0:
        return
    print(msg)


def master_params(optimizer):
    """Generator over the parameters the optimizer actually steps (fp32 masters under O2/O5);
    reference: apex/amp/_amp_state.py:59-69."""
    for group in optimizer.param_groups:
        for p in group["params"]:
            yield p
