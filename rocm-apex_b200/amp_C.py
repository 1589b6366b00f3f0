"""Drop-in `amp_C`: the reference's pybind module (csrc/amp_C_frontend.cpp:147-174), re-implemented
as a thin Python shim over the C ABI of libb200mt.so.

Every function keeps the reference's name and positional signature
`(chunk_size, noop_flag, tensor_lists, *hyperparameters)`, runs asynchronously on torch's current
stream of the tensors' device, and raises RuntimeError with the reference's messages.  Put this
directory first on sys.path and `apex.multi_tensor_apply`, `apex.optimizers`, `apex.amp` (the
reference's Python, or the mirror shipped next to this file) call straight into the sm_100a
kernels.  There is no CPU path.
"""
import torch

import _b200mt as _mt

prepare = _mt.prepare     # extension: amp_C.prepare(tensor_lists) -> a handle every function below accepts as `tensor_lists`

__all__ = [
    "multi_tensor_scale", "multi_tensor_axpby", "multi_tensor_l2norm", "multi_tensor_l2norm_mp",
    "multi_tensor_l2norm_scale", "multi_tensor_adam", "multi_tensor_sgd", "multi_tensor_lamb",
    "multi_tensor_lamb_mp", "multi_tensor_lamb_stage1_cuda", "multi_tensor_lamb_stage2_cuda",
    "multi_tensor_adagrad", "multi_tensor_novograd",
]

_lib = _mt.lib


def multi_tensor_scale(chunk_size, noop_flag, tensor_lists, scale):
    """out = in * scale, overflow flag raised on non-finite input (csrc/amp_C_frontend.cpp:3-7)."""
    L = _mt.pack(tensor_lists, 2)
    _mt.check(_lib.b200mt_scale(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), float(scale),
                                _mt.stream_ptr(L.device)))


def multi_tensor_axpby(chunk_size, noop_flag, tensor_lists, a, b, arg_to_check):
    """out = a*x + b*y (csrc/amp_C_frontend.cpp:24-30)."""
    L = _mt.pack(tensor_lists, 3)
    _mt.check(_lib.b200mt_axpby(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), float(a), float(b),
                                int(arg_to_check), _mt.stream_ptr(L.device)))


def _l2norm(chunk_size, noop_flag, tensor_lists, per_tensor, flags):
    L = _mt.pack(tensor_lists, 1)
    dev = torch.device("cuda", L.device)
    ret = torch.empty(1, dtype=torch.float32, device=dev)
    per = torch.empty(L.ntensors if per_tensor else 0, dtype=torch.float32, device=dev)
    _mt.check(_lib.b200mt_l2norm(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), ret.data_ptr(),
                                 per.data_ptr() if per_tensor else None, flags, _mt.stream_ptr(L.device)))
    return ret, per


def multi_tensor_l2norm(chunk_size, noop_flag, tensor_lists, per_tensor_python=None):
    """-> (norm[1], per_tensor_norm[T] or empty[0]) (csrc/multi_tensor_l2norm_kernel.cu:301-364)."""
    return _l2norm(chunk_size, noop_flag, tensor_lists, bool(per_tensor_python), 0)


def multi_tensor_l2norm_mp(chunk_size, noop_flag, tensor_lists, per_tensor_python=None):
    """As multi_tensor_l2norm but a no-op when the flag is already set
    (csrc/multi_tensor_l2norm_kernel_mp.cu:153-216); outputs are then left uninitialised."""
    return _l2norm(chunk_size, noop_flag, tensor_lists, bool(per_tensor_python), _mt.NORM_NOOP_GATED)


def multi_tensor_l2norm_scale(chunk_size, noop_flag, tensor_lists, scale, per_tensor_python=None):
    """Norm of `in` plus out = in*scale in one pass (csrc/multi_tensor_l2norm_scale_kernel.cu:258-324)."""
    L = _mt.pack(tensor_lists, 2)
    dev = torch.device("cuda", L.device)
    per_tensor = bool(per_tensor_python)
    ret = torch.empty(1, dtype=torch.float32, device=dev)
    per = torch.empty(L.ntensors if per_tensor else 0, dtype=torch.float32, device=dev)
    _mt.check(_lib.b200mt_l2norm_scale(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), float(scale),
                                       ret.data_ptr(), per.data_ptr() if per_tensor else None,
                                       _mt.stream_ptr(L.device)))
    return ret, per


def multi_tensor_adam(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, epsilon, step, mode,
                      bias_correction, weight_decay):
    """[g, p, m, v] Adam (mode 0) / AdamW (mode 1) (csrc/multi_tensor_adam.cu:129-171).
    Extension: a 5th list [.., p_copy] and 16-bit g with fp32 p/m/v run as one fused pass."""
    L = _mt.pack(tensor_lists, (4, 5))
    _mt.check(_lib.b200mt_adam(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), float(lr), float(beta1),
                               float(beta2), float(epsilon), int(step), int(mode), int(bias_correction),
                               float(weight_decay), 1.0, _mt.stream_ptr(L.device)))


def multi_tensor_sgd(chunk_size, noop_flag, tensor_lists, wd, momentum, dampening, lr, nesterov, first_run,
                     wd_after_momentum, scale):
    """[g, p, mom(, p_model_fp16)] (csrc/multi_tensor_sgd_kernel.cu:141-322)."""
    if len(tensor_lists) == 4:
        for t in tensor_lists[3]:
            if t.dtype != torch.float16:
                raise RuntimeError("Additional output tensors should always be fp16.")
    L = _mt.pack(tensor_lists, (3, 4))
    _mt.check(_lib.b200mt_sgd(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag, L.device), float(wd),
                              float(momentum), float(dampening), float(lr), int(bool(nesterov)),
                              int(bool(first_run)), int(bool(wd_after_momentum)), float(scale),
                              _mt.stream_ptr(L.device)))


def multi_tensor_lamb(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, epsilon, step, bias_correction,
                      weight_decay, grad_averaging, mode, global_grad_norm, max_grad_norm, use_nvlamb_python=None):
    """[g, p, m, v] LAMB; g is overwritten with the update (csrc/multi_tensor_lamb.cu:332-413)."""
    L = _mt.pack(tensor_lists, 4)
    _mt.check(_lib.b200mt_lamb(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), float(lr), float(beta1),
                               float(beta2), float(epsilon), int(step), int(bias_correction),
                               float(weight_decay), int(grad_averaging), int(mode),
                               _mt.dev_ptr(global_grad_norm, torch.float32, "global_grad_norm"),
                               float(max_grad_norm), int(bool(use_nvlamb_python)), _mt.stream_ptr(L.device)))


def multi_tensor_lamb_mp(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, epsilon, step, bias_correction,
                         weight_decay, grad_averaging, mode, global_grad_norm, max_grad_norm,
                         use_nvlamb_python, found_inf, inv_scale):
    """Sync-free mixed-precision LAMB: lr/step/norms/scales are device tensors
    (csrc/multi_tensor_lamb_mp.cu:367-496)."""
    L = _mt.pack(tensor_lists, (4, 5))
    _mt.check(_lib.b200mt_lamb_mp(
        L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), _mt.dev_ptr(lr, torch.float32, "lr"),
        float(beta1), float(beta2), float(epsilon), _mt.dev_ptr(step, torch.int32, "step"),
        int(bias_correction), float(weight_decay), int(grad_averaging), int(mode),
        _mt.dev_ptr(global_grad_norm, torch.float32, "global_grad_norm"),
        _mt.dev_ptr(max_grad_norm, torch.float32, "max_grad_norm"), int(bool(use_nvlamb_python)),
        _mt.dev_ptr(found_inf, torch.float32, "found_inf"), _mt.dev_ptr(inv_scale, torch.float32, "inv_scale"),
        _mt.stream_ptr(L.device)))


def multi_tensor_lamb_stage1_cuda(chunk_size, noop_flag, tensor_lists, per_tensor_decay, step, beta1, beta2,
                                  epsilon, global_grad_norm, max_global_grad_norm):
    """[g, p, m, v, update]; global_grad_norm is dereferenced on the host, as in the reference
    (csrc/multi_tensor_lamb_stage_1.cu:126-127), so it must be a CPU tensor."""
    if global_grad_norm.is_cuda:
        raise RuntimeError("multi_tensor_lamb_stage1_cuda reads global_grad_norm on the host: pass a CPU tensor")
    gn = float(global_grad_norm.reshape(-1)[0])
    clip = gn / max_global_grad_norm if gn > max_global_grad_norm else 1.0
    L = _mt.pack(tensor_lists, 5)
    _mt.check(_lib.b200mt_lamb_stage1(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag),
                                      _mt.dev_ptr(per_tensor_decay, torch.float32, "per_tensor_decay"),
                                      int(step), float(beta1), float(beta2), float(epsilon), float(clip),
                                      _mt.stream_ptr(L.device)))


def multi_tensor_lamb_stage2_cuda(chunk_size, noop_flag, tensor_lists, per_tensor_param_norm,
                                  per_tensor_update_norm, lr, weight_decay, use_nvlamb_python=None):
    """[p, update] trust-ratio apply (csrc/multi_tensor_lamb_stage_2.cu:94-125)."""
    L = _mt.pack(tensor_lists, 2)
    _mt.check(_lib.b200mt_lamb_stage2(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag),
                                      _mt.dev_ptr(per_tensor_param_norm, torch.float32, "per_tensor_param_norm"),
                                      _mt.dev_ptr(per_tensor_update_norm, torch.float32, "per_tensor_update_norm"),
                                      float(lr), float(weight_decay), int(bool(use_nvlamb_python)),
                                      _mt.stream_ptr(L.device)))


def multi_tensor_adagrad(chunk_size, noop_flag, tensor_lists, lr, epsilon, mode, weight_decay):
    """[g, p, h] Adagrad, mode 0 L2 / 1 decoupled decay (csrc/multi_tensor_adagrad.cu:85-100)."""
    L = _mt.pack(tensor_lists, 3)
    _mt.check(_lib.b200mt_adagrad(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), float(lr), float(epsilon),
                                  int(mode), float(weight_decay), _mt.stream_ptr(L.device)))


def multi_tensor_novograd(chunk_size, noop_flag, tensor_lists, grad_norms, lr, beta1, beta2, epsilon, step,
                          bias_correction, weight_decay, grad_averaging, moment_mode, norm_type):
    """[g, p, m] NovoGrad; grad_norms (fp32[T], device) is blended with this step's per-tensor gradient
    norm in place, then used as the second moment (csrc/multi_tensor_novograd.cu:120-188)."""
    L = _mt.pack(tensor_lists, 3)
    if grad_norms.numel() != L.ntensors:
        raise RuntimeError("grad_norms must hold one value per tensor")
    _mt.check(_lib.b200mt_novograd(L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag, L.device),
                                   _mt.dev_ptr(grad_norms, torch.float32, "grad_norms"), float(lr), float(beta1),
                                   float(beta2), float(epsilon), int(step), int(bias_correction), float(weight_decay),
                                   int(grad_averaging), int(moment_mode), int(norm_type), _mt.stream_ptr(L.device)))
