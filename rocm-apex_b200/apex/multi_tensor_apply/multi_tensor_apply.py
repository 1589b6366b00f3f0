"""The launch shim of the reference (apex/multi_tensor_apply/multi_tensor_apply.py:3-34): a callable
that forwards (chunk_size, noop_flag, tensor_lists, *args) to an `amp_C` op.  `available` reflects
whether `amp_C` (here: the shim over libb200mt.so) could be imported."""


class MultiTensorApply(object):
    available = False
    warned = False

    def __init__(self, chunk_size):
        try:
            import amp_C  # noqa: F401
            MultiTensorApply.available = True
            self.chunk_size = chunk_size
        except (ImportError, RuntimeError, OSError) as err:
            MultiTensorApply.available = False
            MultiTensorApply.import_err = err

    def check_avail(self):
        if not MultiTensorApply.available:
            raise RuntimeError(
                "Attempted to call MultiTensorApply method, but MultiTensorApply is not available, "
                "possibly because libb200mt.so is not built.  Original import error message:",
                MultiTensorApply.import_err)

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        self.check_avail()
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
