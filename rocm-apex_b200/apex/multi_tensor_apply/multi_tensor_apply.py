"""The launch shim every fused optimizer and amp's scaler go through (reference:
apex/multi_tensor_apply/multi_tensor_apply.py:3-34): `multi_tensor_applier(op, noop_flag, tensor_lists, *args)`
calls `op(chunk_size, noop_flag, tensor_lists, *args)`.  Whether an `amp_C` exists -- here the ctypes shim over
libb200mt.so -- is decided once, kept on the class as the reference keeps it (`available`, `import_err`), and
reported with the reference's message when somebody calls through a shim that has nothing behind it."""


def _why_amp_C_is_missing():
    """None when `import amp_C` works; otherwise the exception (a missing or unloadable libb200mt.so surfaces as
    RuntimeError / OSError from the binding, not only as ImportError)."""
    try:
        import amp_C  # noqa: F401
    except (ImportError, RuntimeError, OSError) as exc:
        return exc
    return None


class MultiTensorApply(object):
    available = False
    warned = False
    import_err = None

    def __init__(self, chunk_size):
        self.chunk_size = chunk_size
        MultiTensorApply.import_err = _why_amp_C_is_missing()
        MultiTensorApply.available = MultiTensorApply.import_err is None

    def check_avail(self):
        if MultiTensorApply.available:
            return
        raise RuntimeError("Attempted to call MultiTensorApply method, but MultiTensorApply is not available, "
                           "possibly because libb200mt.so is not built.  Original import error message:",
                           MultiTensorApply.import_err)

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        self.check_avail()
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
