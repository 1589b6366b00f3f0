"""Helpers shared by the fused optimizers of this mirror."""
import torch

from ..multi_tensor_apply import multi_tensor_applier

SUPPORTED = (torch.float32, torch.float16, torch.bfloat16)


def require_extension(name):
    if not multi_tensor_applier.available:
        raise RuntimeError("apex.optimizers.%s requires cuda extensions" % name)
    import amp_C
    return amp_C


def overflow_buf(device=None):
    return torch.zeros(1, dtype=torch.int32, device=device if device is not None else "cuda")


def by_dtype(entries, key=1):
    """Groups (grad, param, ...) tuples by the dtype of member `key` (the parameter), fp32 first.  One launch
    per dtype: the reference lumps fp16 and bf16 into one list and dispatches on its first tensor
    (apex/optimizers/fused_adam.py:134-138), which mistypes mixed lists; keeping them apart does not."""
    out = {}
    for e in entries:
        out.setdefault(e[key].dtype, []).append(e)
    return [out[d] for d in SUPPORTED if d in out]
