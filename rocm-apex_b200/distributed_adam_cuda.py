"""Drop-in `distributed_adam_cuda` (apex/contrib/csrc/optimizers/multi_tensor_distopt_adam.cpp:3-20):
the shard-local fused AdamW step of the ZeRO-style DistributedFusedAdam."""
import torch

import _b200mt as _mt

_lib = _mt.lib


def multi_tensor_fused_adam(chunk_size, noop_flag, tensor_lists, per_tensor_beta1, per_tensor_beta2,
                            per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale,
                            step, mode):
    """tensor_lists = [p32, m32, v32, g(, p_copy)]; per-tensor hyper-parameters are device arrays
    (multi_tensor_distopt_adam_kernel.cu:169-228).  mode 0: eps inside sqrt, 1: outside."""
    if len(tensor_lists) not in (4, 5):
        raise RuntimeError("expected tensor lists of size 4 or 5")
    L = _mt.pack(tensor_lists, (4, 5))
    _mt.check(_lib.b200mt_dist_adam(
        L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag),
        _mt.dev_ptr(per_tensor_beta1, torch.float32, "per_tensor_beta1"),
        _mt.dev_ptr(per_tensor_beta2, torch.float32, "per_tensor_beta2"),
        _mt.dev_ptr(per_tensor_bias_correction, torch.int32, "per_tensor_bias_correction"),
        _mt.dev_ptr(per_tensor_eps, torch.float32, "per_tensor_eps"),
        _mt.dev_ptr(per_tensor_weight_decay, torch.float32, "per_tensor_weight_decay"),
        float(lr), float(grad_scale), int(step), int(mode), _mt.stream_ptr(L.device)))


def multi_tensor_fused_adam_nvls(chunk_size, noop_flag, tensor_lists, per_tensor_beta1, per_tensor_beta2,
                                 per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale,
                                 step, mode):
    """Extension (no reference counterpart): same arguments as multi_tensor_fused_adam, but lists 3 and 4
    ([.., g, p_copy]) are `_b200mt.RawBuffer`s addressing NVSwitch multicast memory, so that the gradient
    reduce-scatter and the parameter all-gather happen inside this one kernel (include/b200mt.h,
    b200mt_dist_adam_nvls).  The caller provides the cross-GPU barriers around it."""
    if len(tensor_lists) != 5:
        raise RuntimeError("expected tensor lists of size 5")
    L = _mt.pack(tensor_lists, 5)
    _mt.check(_lib.b200mt_dist_adam_nvls(
        L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag),
        _mt.dev_ptr(per_tensor_beta1, torch.float32, "per_tensor_beta1"),
        _mt.dev_ptr(per_tensor_beta2, torch.float32, "per_tensor_beta2"),
        _mt.dev_ptr(per_tensor_bias_correction, torch.int32, "per_tensor_bias_correction"),
        _mt.dev_ptr(per_tensor_eps, torch.float32, "per_tensor_eps"),
        _mt.dev_ptr(per_tensor_weight_decay, torch.float32, "per_tensor_weight_decay"),
        float(lr), float(grad_scale), int(step), int(mode), _mt.stream_ptr(L.device)))


def multi_tensor_fused_adam_p2p(chunk_size, noop_flag, tensor_lists, per_tensor_beta1, per_tensor_beta2,
                                per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale,
                                step, mode, grad_peer_delta, copy_peer_delta):
    """Extension (no reference counterpart): the sharded step of TWO ranks as one kernel over peer-mapped memory
    (include/b200mt.h, b200mt_dist_adam_p2p).  Lists 3 and 4 are this rank's views of symmetric buffers; the deltas are the
    byte distances to the peer's copies.  The caller provides the cross-GPU barriers around it."""
    if len(tensor_lists) != 5:
        raise RuntimeError("expected tensor lists of size 5")
    L = _mt.pack(tensor_lists, 5)
    _mt.check(_lib.b200mt_dist_adam_p2p(
        L.ref(), int(chunk_size), _mt.noop_ptr(noop_flag),
        _mt.dev_ptr(per_tensor_beta1, torch.float32, "per_tensor_beta1"),
        _mt.dev_ptr(per_tensor_beta2, torch.float32, "per_tensor_beta2"),
        _mt.dev_ptr(per_tensor_bias_correction, torch.int32, "per_tensor_bias_correction"),
        _mt.dev_ptr(per_tensor_eps, torch.float32, "per_tensor_eps"),
        _mt.dev_ptr(per_tensor_weight_decay, torch.float32, "per_tensor_weight_decay"),
        float(lr), float(grad_scale), int(step), int(mode), int(grad_peer_delta), int(copy_peer_delta),
        _mt.stream_ptr(L.device)))
