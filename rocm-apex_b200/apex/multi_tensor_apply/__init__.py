from .multi_tensor_apply import MultiTensorApply

# the reference's global chunk size (apex/multi_tensor_apply/__init__.py:3)
multi_tensor_applier = MultiTensorApply(2048 * 32)
