"""Drop-in `distributed_lamb_cuda` (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:32-36): the two
kernels of the sharded LAMB step.  Same names and positional signatures as the reference; `update_sumsq` is
an optional extension of compute_update_term (see include/b200mt.h)."""
import torch

import _b200mt as _mt

_lib = _mt.lib


def multi_tensor_lamb_compute_update_term(chunk_size, noop_flag, tensor_lists, per_tensor_beta1, per_tensor_beta2,
                                          per_tensor_beta3, per_tensor_bias_correction, step, per_tensor_epsilon, mode,
                                          per_tensor_decay, global_scale, global_grad_norm, max_grad_norm,
                                          update_sumsq=None):
    """tensor_lists = [g, p, m, v, u]; step (int32), global_scale and global_grad_norm (fp32) are device tensors."""
    L = _mt.pack(tensor_lists, 5)
    f32, i32 = torch.float32, torch.int32
    _mt.check(_lib.b200mt_dist_lamb_stage1(
        L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), _mt.dev_ptr(per_tensor_beta1, f32, "per_tensor_beta1"),
        _mt.dev_ptr(per_tensor_beta2, f32, "per_tensor_beta2"), _mt.dev_ptr(per_tensor_beta3, f32, "per_tensor_beta3"),
        _mt.dev_ptr(per_tensor_bias_correction, i32, "per_tensor_bias_correction"), _mt.dev_ptr(step, i32, "step"),
        _mt.dev_ptr(per_tensor_epsilon, f32, "per_tensor_epsilon"), int(mode), _mt.dev_ptr(per_tensor_decay, f32, "per_tensor_decay"),
        _mt.dev_ptr(global_scale, f32, "global_scale"), _mt.dev_ptr(global_grad_norm, f32, "global_grad_norm"),
        float(max_grad_norm), _mt.dev_ptr(update_sumsq, f32, "update_sumsq") if update_sumsq is not None else None,
        _mt.stream_ptr(L.device)))


def multi_tensor_lamb_update_weights(chunk_size, noop_flag, tensor_lists, per_tensor_param_norm, per_tensor_update_norm,
                                     update_norm_offset, learning_rate, per_tensor_decay, global_grad_norm, use_nvlamb):
    """tensor_lists = [u, p, p_copy]; update_norm_offset (int64, device) maps a fragment to its parameter's update norm."""
    L = _mt.pack(tensor_lists, 3)
    f32 = torch.float32
    _mt.check(_lib.b200mt_dist_lamb_stage2(
        L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag), _mt.dev_ptr(per_tensor_param_norm, f32, "per_tensor_param_norm"),
        _mt.dev_ptr(per_tensor_update_norm, f32, "per_tensor_update_norm"),
        _mt.dev_ptr(update_norm_offset, torch.int64, "update_norm_offset"), _mt.dev_ptr(learning_rate, f32, "learning_rate"),
        _mt.dev_ptr(per_tensor_decay, f32, "per_tensor_decay"), _mt.dev_ptr(global_grad_norm, f32, "global_grad_norm"),
        int(bool(use_nvlamb)), _mt.stream_ptr(L.device)))
