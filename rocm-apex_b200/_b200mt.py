"""ctypes binding of libb200mt.so (declared in include/b200mt.h) plus the tensor-list packer.

This is the only place where torch tensors are turned into the raw (pointer, numel, dtype) arrays
the C ABI takes.  The checks performed while packing, and their messages, are the ones the
reference's launch harness performs (csrc/multi_tensor_apply.cuh:53-70) so that callers and tests
written against the reference see the same errors.

There is no CPU path: importing this module without the built library raises immediately.
"""
import ctypes
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200MT_LIB: load another build of the same library (tuning experiments with different compile-time
# vector widths); the default is the in-tree build next to this file.
_LIB_PATH = os.environ.get("B200MT_LIB") or os.path.join(_HERE, "libb200mt.so")

F32, F16, BF16, F64, U8 = 0, 1, 2, 3, 4
MAX_DEPTH = 5
NORM_NOOP_GATED = 1

_DTYPE_CODE = {
    torch.float32: F32,
    torch.float16: F16,
    torch.bfloat16: BF16,
    torch.float64: F64,
    torch.uint8: U8,
}


class Lists(ctypes.Structure):
    _fields_ = [
        ("ptrs", ctypes.c_void_p),
        ("numel", ctypes.c_void_p),
        ("ntensors", ctypes.c_int32),
        ("depth", ctypes.c_int32),
        ("dtype", ctypes.c_int32 * MAX_DEPTH),
        ("device", ctypes.c_int32),
        ("prepared", ctypes.c_void_p),
    ]


def _load():
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(
            "libb200mt.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C rocm-apex_b200`. There is no CPU or PyTorch fallback for this path." % _LIB_PATH)
    lib = ctypes.CDLL(_LIB_PATH)
    c_int, c_float, vp = ctypes.c_int, ctypes.c_float, ctypes.c_void_p
    LP = ctypes.POINTER(Lists)
    sigs = {
        "b200mt_last_error": (ctypes.c_char_p, []),
        "b200mt_version": (c_int, []),
        "b200mt_launch_count": (ctypes.c_int64, []),
        "b200mt_table_uploads": (ctypes.c_int64, []),
        "b200mt_table_hits": (ctypes.c_int64, []),
        "b200mt_set_tuning": (c_int, [c_int, c_int]),
        "b200mt_shutdown": (c_int, []),
        "b200mt_lists_prepare": (c_int, [LP, ctypes.POINTER(ctypes.c_void_p)]),
        "b200mt_lists_release": (c_int, [vp]),
        "b200mt_trace_begin": (c_int, [c_int]),
        "b200mt_trace_end": (c_int, [vp, c_int, vp]),
        "b200mt_trace_name": (ctypes.c_char_p, [c_int]),
        "b200mt_scale": (c_int, [LP, c_int, vp, c_float, vp]),
        "b200mt_axpby": (c_int, [LP, c_int, vp, c_float, c_float, c_int, vp]),
        "b200mt_l2norm": (c_int, [LP, c_int, vp, vp, vp, c_int, vp]),
        "b200mt_l2norm_scale": (c_int, [LP, c_int, vp, c_float, vp, vp, vp]),
        "b200mt_adam": (c_int, [LP, c_int, vp, c_float, c_float, c_float, c_float, c_int, c_int, c_int,
                                c_float, c_float, vp]),
        "b200mt_sgd": (c_int, [LP, c_int, vp, c_float, c_float, c_float, c_float, c_int, c_int, c_int,
                               c_float, vp]),
        "b200mt_lamb": (c_int, [LP, c_int, vp, c_float, c_float, c_float, c_float, c_int, c_int, c_float,
                                c_int, c_int, vp, c_float, c_int, vp]),
        "b200mt_lamb_mp": (c_int, [LP, c_int, vp, vp, c_float, c_float, c_float, vp, c_int, c_float, c_int,
                                   c_int, vp, vp, c_int, vp, vp, vp]),
        "b200mt_lamb_stage1": (c_int, [LP, c_int, vp, vp, c_int, c_float, c_float, c_float, c_float, vp]),
        "b200mt_lamb_stage2": (c_int, [LP, c_int, vp, vp, vp, c_float, c_float, c_int, vp]),
        "b200mt_dist_adam": (c_int, [LP, c_int, vp, vp, vp, vp, vp, vp, c_float, c_float, c_int, c_int, vp]),
        "b200mt_dist_adam_nvls": (c_int, [LP, c_int, vp, vp, vp, vp, vp, vp, c_float, c_float, c_int, c_int, vp]),
        "b200mt_dist_adam_p2p": (c_int, [LP, c_int, vp, vp, vp, vp, vp, vp, c_float, c_float, c_int, c_int, ctypes.c_int64,
                                         ctypes.c_int64, vp]),
        "b200mt_adagrad": (c_int, [LP, c_int, vp, c_float, c_float, c_int, c_float, vp]),
        "b200mt_novograd": (c_int, [LP, c_int, vp, vp, c_float, c_float, c_float, c_float, c_int, c_int, c_float, c_int,
                                    c_int, c_int, vp]),
        "b200mt_dist_lamb_stage1": (c_int, [LP, c_int, vp, vp, vp, vp, vp, vp, vp, c_int, vp, vp, vp, c_float, vp, vp]),
        "b200mt_dist_lamb_stage2": (c_int, [LP, c_int, vp, vp, vp, vp, vp, vp, vp, c_int, vp]),
        "b200mt_nvls_reduce_scatter": (c_int, [LP, c_int, vp]),
        "b200mt_nvls_all_gather": (c_int, [LP, c_int, vp]),
        "b200mt_contrib_adam": (c_int, [LP, c_int, vp, c_float, c_float, c_float, c_float, c_float, c_int,
                                        c_int, c_int, c_float, vp]),
        "b200mt_reversible_adam": (c_int, [LP, c_int, c_float, c_float, c_float, c_float, c_float, c_int, c_int, c_int,
                                           c_float, vp]),
        "b200mt_maybe_adam_undo": (c_int, [LP, c_int, vp, c_float, c_float, c_float, c_float, c_float, c_int, c_int,
                                           c_int, c_float, vp]),
        "b200mt_maybe_cast": (c_int, [LP, c_int, vp, vp]),
        "b200mt_strided_check_finite": (c_int, [vp, vp, c_int, ctypes.c_int64, c_int, c_int, c_int, vp]),
        "b200mt_l2norm_flat_noalloc": (c_int, [vp, c_int, ctypes.c_int64, c_int, vp, vp, c_int, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)  # AttributeError here = header and library out of sync: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib, tuple(sigs)


lib, EXPORTS = _load()

try:   # optional native list walker (build_packer.py); the Python walker below is its exact equivalent
    from _b200mt_pack import pack_lists as _fast_pack
except ImportError:
    _fast_pack = None
try:
    from _b200mt_pack import check_rows
except ImportError:
    check_rows = None
if os.environ.get("B200MT_PY_PACK"):
    _fast_pack = check_rows = None


class trace(object):
    """with trace() as t: ...; t.launches -> [(kernel name, ms)] for every kernel the library launched inside the block
    (b200mt_trace_begin / _end: events around each launch; PDL off while tracing)."""

    def __init__(self, max_launches=4096):
        self.max, self.launches = max_launches, []

    def __enter__(self):
        check(lib.b200mt_trace_begin(self.max))
        return self

    def __exit__(self, *exc):
        ms = (ctypes.c_float * self.max)()
        n = ctypes.c_int(0)
        check(lib.b200mt_trace_end(ms, self.max, ctypes.byref(n)))
        self.launches = [(lib.b200mt_trace_name(i).decode(), float(ms[i])) for i in range(min(n.value, self.max))]
        return False


def check(rc):
    """Turn a non-zero C return code into the RuntimeError the reference's TORCH_CHECK would raise."""
    if rc != 0:
        raise RuntimeError(lib.b200mt_last_error().decode("utf-8", "replace"))


def dtype_code(t):
    try:
        return _DTYPE_CODE[t.dtype]
    except KeyError:
        raise RuntimeError("multi_tensor_apply not implemented for '%s'" % str(t.dtype).replace("torch.", ""))


class RawBuffer(object):
    """A device address range that is not a torch tensor -- e.g. an address inside a CUDA multicast
    object, which no tensor may alias.  Quacks enough like a tensor for pack()."""
    is_sparse = False

    def __init__(self, ptr, numel, dtype, device):
        self._ptr, self._numel, self.dtype, self.device = int(ptr), int(numel), dtype, device

    def data_ptr(self):
        return self._ptr

    def numel(self):
        return self._numel

    def dim(self):
        return 1

    def is_contiguous(self, memory_format=None):
        return True


class Packed(object):
    """A tensor-list set in C-ABI form.  Keeps the numpy arrays alive for the duration of the call."""
    __slots__ = ("struct", "ptrs", "numel", "device", "ntensors", "depth", "handle", "keep", "__weakref__")

    def ref(self):
        return ctypes.byref(self.struct)


class Prepared(Packed):
    """A list set prepared once (b200mt_lists_prepare): every drop-in entry point accepts it in place of
    `tensor_lists` and then skips the per-call walk of the lists, the table image and the compare -- what
    is left of a call is the launch.  The caller promises that the tensors stay at their addresses and sizes for the
    life of the object (it keeps the tensor objects alive, so freed storage cannot be reused under it; re-pointing a
    tensor with `.data =` / `set_()` / `resize_()` is the one thing that breaks the promise).  release() or garbage
    collection frees the device tables."""
    __slots__ = ()

    def __len__(self):             # quacks like the list of lists it was made from
        return self.depth

    def __getitem__(self, d):
        return self.keep[d]

    def release(self):
        h, self.handle = getattr(self, "handle", None), None
        if h:
            self.struct.prepared = None
            lib.b200mt_lists_release(h)

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


def prepare(tensor_lists, keep=True):
    """-> Prepared: validate and pack `tensor_lists` once, keep a resident descriptor table for them.
    keep=False: the handle does not hold the tensors; the caller re-validates the addresses (q.ptrs) before every use."""
    p = pack(tensor_lists)
    q = Prepared()
    for name in ("struct", "ptrs", "numel", "device", "ntensors", "depth"):
        setattr(q, name, getattr(p, name))
    q.keep = [list(l) for l in tensor_lists] if keep else None   # the tensors stay alive as long as the handle does
    h = ctypes.c_void_p()
    check(lib.b200mt_lists_prepare(q.ref(), ctypes.byref(h)))
    q.handle = h.value
    q.struct.prepared = h.value
    return q


def pack(tensor_lists, depth=None):
    """Validate `tensor_lists` like csrc/multi_tensor_apply.cuh:53-70 and build the C descriptor.
    A `Prepared` list set passes straight through (only its depth is checked)."""
    if isinstance(tensor_lists, Packed):
        if depth is not None:
            ok = (tensor_lists.depth in depth) if isinstance(depth, (tuple, list)) else (tensor_lists.depth == depth)
            if not ok:
                raise RuntimeError("tensor_lists.size() != depth")
        return tensor_lists
    nlists = len(tensor_lists)
    if depth is not None:
        ok = (nlists in depth) if isinstance(depth, (tuple, list)) else (nlists == depth)
        if not ok:
            raise RuntimeError("tensor_lists.size() != depth")
    if nlists < 1 or nlists > MAX_DEPTH:
        raise RuntimeError("tensor_lists.size() != depth")
    first = tensor_lists[0]
    ntensors = len(first)
    if ntensors <= 0:
        raise RuntimeError("tensor_lists[0].size() is not > 0")
    if _fast_pack is not None:
        # native walk of the lists (csrc/packer.cpp): same checks, same messages, ~50x less host time
        ptrs = np.empty(nlists * ntensors, dtype=np.uint64)
        numel = np.empty(ntensors, dtype=np.int64)
        r = _fast_pack(tensor_lists, ptrs.ctypes.data, numel.ctypes.data, nlists * ntensors)
        if r is not None:
            _, _, codes, dev_index = r
            p = Packed()
            p.ptrs, p.numel, p.ntensors, p.depth, p.device = ptrs, numel, ntensors, nlists, dev_index
            s = Lists()
            s.ptrs, s.numel, s.ntensors, s.depth, s.device = ptrs.ctypes.data, numel.ctypes.data, ntensors, nlists, dev_index
            for d, c in enumerate(codes):
                s.dtype[d] = c
            p.struct = s
            return p
        # r is None: some entry is not a torch tensor (e.g. a RawBuffer) -> generic path below
    ref_device = first[0].device
    if ref_device.type != "cuda":
        raise RuntimeError("expected input to be on cuda")
    ptrs = np.empty(nlists * ntensors, dtype=np.uint64)
    numel = np.empty(ntensors, dtype=np.int64)
    for t in range(ntensors):
        numel[t] = first[t].numel()
    k = 0
    for lst in tensor_lists:
        if len(lst) != ntensors:
            raise RuntimeError("Size mismatch among tensor lists")
        dt = lst[0].dtype
        for t, x in enumerate(lst):
            if x.is_sparse:
                raise RuntimeError("sparse tensors are not supported on this path")
            if not (x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))
                    or (x.dim() == 5 and x.is_contiguous(memory_format=torch.channels_last_3d))):
                raise RuntimeError("A tensor was not contiguous.")
            if x.device != ref_device:
                raise RuntimeError("A tensor was not on the same device as the first tensor")
            if x.numel() != numel[t]:
                raise RuntimeError("Size mismatch")
            if x.dtype != dt:
                raise RuntimeError("tensors of one list must share a dtype (tensor %d)" % t)
            ptrs[k] = x.data_ptr()
            k += 1
    p = Packed()
    p.ptrs, p.numel = ptrs, numel
    p.ntensors, p.depth = ntensors, nlists
    p.device = ref_device.index if ref_device.index is not None else torch.cuda.current_device()
    s = Lists()
    s.ptrs = ptrs.ctypes.data
    s.numel = numel.ctypes.data
    s.ntensors = ntensors
    s.depth = nlists
    for d, lst in enumerate(tensor_lists):
        s.dtype[d] = dtype_code(lst[0])
    s.device = p.device
    p.struct = s
    return p


def stream_ptr(device_index):
    """cudaStream_t of torch's current stream on the tensors' device (the reference launches on
    at::cuda::getCurrentCUDAStream() under a device guard, csrc/multi_tensor_apply.cuh:78-79)."""
    return torch.cuda.current_stream(device_index).cuda_stream


def dev_ptr(t, dtype=None, what="tensor"):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("expected %s to be on cuda" % what)
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError("expected %s to have dtype %s, got %s" % (what, dtype, t.dtype))
    return t.data_ptr()


def noop_ptr(noop_flag, device_index=None):
    if not noop_flag.is_cuda:
        raise RuntimeError("expected noop flag to be on cuda")
    if noop_flag.dtype != torch.int32:
        raise RuntimeError("expected noop flag to be an int32 tensor")
    if device_index is not None and noop_flag.device.index != device_index:
        raise RuntimeError("expected noop flag to be on the same device as tensors")
    return noop_flag.data_ptr()
