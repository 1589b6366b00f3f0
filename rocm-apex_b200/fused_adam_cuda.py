"""Drop-in `fused_adam_cuda` (apex/contrib/csrc/optimizers/fused_adam_cuda.cpp:78-86), all eight
entry points: adam / adam_mt, reversible_adam, maybe_adam_undo, maybe_cast / maybe_cast_mt, strided_check_finite."""
import torch

import _b200mt as _mt

_lib = _mt.lib


def adam_mt(chunk_size, overflow_flag, tensor_lists, lr, beta1, beta2, eps, grad_scale, step, mode,
            bias_correction, decay):
    """tensor_lists = [p, m, v, g(, p_copy)], step_size = lr*sqrt(bc2)/bc1 form
    (fused_adam_cuda_kernel.cu:283-387)."""
    if len(tensor_lists) not in (4, 5):
        raise RuntimeError("expected tensor lists of size 4 or 5")
    L = _mt.pack(tensor_lists, (4, 5))
    _mt.check(_lib.b200mt_contrib_adam(L.ref(), int(chunk_size), _mt.noop_ptr(overflow_flag), float(lr),
                                       float(beta1), float(beta2), float(eps), float(grad_scale), int(step),
                                       int(mode), int(bias_correction), float(decay), _mt.stream_ptr(L.device)))


def maybe_cast_mt(chunk_size, overflow_flag, tensor_lists):
    """tensor_lists = [p_in, p_out]; out = (out dtype)in unless the flag is set (:939-960)."""
    if len(tensor_lists) != 2:
        raise RuntimeError("expected tensor lists of size 2")
    L = _mt.pack(tensor_lists, 2)
    _mt.check(_lib.b200mt_maybe_cast(L.ref(), int(chunk_size), _mt.noop_ptr(overflow_flag),
                                     _mt.stream_ptr(L.device)))


def strided_check_finite(overflow_flag, p_copy, stride, clear_overflow_first):
    """Tests every stride-th element of p_copy; optionally clears the flag first (:785-812)."""
    if not p_copy.is_cuda or not p_copy.is_contiguous():
        raise RuntimeError("p_copy must be a contiguous CUDA tensor")
    dev = p_copy.device.index
    _mt.check(_lib.b200mt_strided_check_finite(_mt.noop_ptr(overflow_flag), p_copy.data_ptr(),
                                               _mt.dtype_code(p_copy), p_copy.numel(), int(stride),
                                               int(clear_overflow_first), dev, _mt.stream_ptr(dev)))


def adam(p, p_copy, m, v, g, lr, beta1, beta2, eps, grad_scale, step, mode, bias_correction, decay):
    """Flat-tensor form of adam_mt (fused_adam_cuda.cpp:27-38, fused_adam_cuda_kernel.cu:202-281): p_copy may be an
    empty tensor when no low-precision copy is wanted."""
    n = p.numel()
    for name, t in (("m", m), ("v", v), ("g", g)):
        if t.numel() != n:
            raise RuntimeError("number of elements in %s and p tensors should be equal" % name)
    if p_copy.numel() not in (0, n):
        raise RuntimeError("number of elements in p_copy and p tensors should be equal, or p_copy should be empty")
    lists = [[p], [m], [v], [g]] + ([[p_copy]] if p_copy.numel() else [])
    dummy = torch.zeros(1, dtype=torch.int32, device=p.device)
    adam_mt(2048 * 32, dummy, lists, lr, beta1, beta2, eps, grad_scale, step, mode, bias_correction, decay)


def maybe_cast(overflow_flag, p_in, p_out):
    """Flat-tensor form of maybe_cast_mt (fused_adam_cuda.cpp:68-75)."""
    if p_out.numel() != p_in.numel():
        raise RuntimeError("number of elements in p_in and p_out should be equal")
    maybe_cast_mt(2048 * 32, overflow_flag, [[p_in], [p_out]])


def _check_flat(p, m, v, g):
    n = p.numel()
    for name, t in (("m", m), ("v", v), ("g", g)):
        if t.numel() != n:
            raise RuntimeError("number of elements in %s and p tensors should be equal" % name)
    return n


def reversible_adam(p, p_copy, m, v, g, lr, beta1, beta2, eps, grad_scale, step, mode, bias_correction, decay):
    """adam() applied only where g/grad_scale is finite; on overflow the elements concerned stay put and p_copy[0]
    becomes +inf, to be found by strided_check_finite (fused_adam_cuda.cpp:43-55, fused_adam_cuda_kernel.cu:570-653).
    An fp32 gradient writes no p_copy, like the reference (:897-899)."""
    n = _check_flat(p, m, v, g)
    if p_copy.numel() not in (0, n):
        raise RuntimeError("number of elements in p_copy and p tensors should be equal, or p_copy should be empty")
    lists = [[p], [m], [v], [g]] + ([[p_copy]] if (p_copy.numel() and g.dtype != torch.float32) else [])
    L = _mt.pack(lists, (4, 5))
    _mt.check(_lib.b200mt_reversible_adam(L.ref(), 2048 * 32, float(lr), float(beta1), float(beta2), float(eps),
                                          float(grad_scale), int(step), int(mode), int(bias_correction), float(decay),
                                          _mt.stream_ptr(L.device)))


def maybe_adam_undo(overflow_flag, p, m, v, g, lr, beta1, beta2, eps, grad_scale, step, mode, bias_correction, decay):
    """Inverse of reversible_adam's step, applied when the flag is set (or the flag tensor is empty)
    (fused_adam_cuda.cpp:57-67, fused_adam_cuda_kernel.cu:655-732).  Reverts EVERY element; the reference's kernel
    only reaches elements 0 mod 4 (see csrc/ops_reversible.cu)."""
    _check_flat(p, m, v, g)
    L = _mt.pack([[p], [m], [v], [g]], 4)
    flag = _mt.noop_ptr(overflow_flag) if overflow_flag.numel() else None
    _mt.check(_lib.b200mt_maybe_adam_undo(L.ref(), 2048 * 32, flag, float(lr), float(beta1), float(beta2), float(eps),
                                          float(grad_scale), int(step), int(mode), int(bias_correction), float(decay),
                                          _mt.stream_ptr(L.device)))
