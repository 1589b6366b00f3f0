"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the sharded LAMB kernels of DistributedFusedLAMB
(apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu).  Pinned against the reference's own kernels:
oracle/build_ref.sh compiles that file unmodified for sm_100a (with the -D__CUDA_NO_HALF*_ defines torch's CUDAExtension
adds, without which :352 meets an ambiguous operator!=) into oracle/_ref/distributed_lamb_cuda_ref.so, and
tests/test_dist_lamb.py::test_dist_lamb_kernels_match_reference_binary compares this restatement AND the sm_100a kernels
of csrc/ops_dist_lamb.cu with it on identical inputs.
"""
import math

import numpy as np

from .mt_oracle import _to_np, _store, _round_trip, f32


def compute_update_term(noop, lists, b1s, b2s, bias_corrs, step, epss, mode, decays, global_scale, global_grad_norm,
                        max_grad_norm):
    """multi_tensor_distopt_lamb_kernel.cu:126-160,192-236.  lists = [g, p, m, v, u].  Returns the per-fragment sum
    of squares of the stored update (what the reference obtains with a separate l2norm, distributed_fused_lamb.py:686-689)."""
    if noop[0] == 1:
        return None
    gs = f32(global_scale)
    cs = gs
    if max_grad_norm > 0:
        cs = f32(max_grad_norm) / (f32(global_grad_norm) / gs + f32(1e-6))
        cs = gs / min(f32(1.0), cs)
    one = f32(1)
    sums = []
    for k in range(len(lists[0])):
        g, p, m, v = (_to_np(lists[d][k]) for d in range(4))
        b1, b2, eps, decay = f32(b1s[k]), f32(b2s[k]), f32(epss[k]), f32(decays[k])
        b3 = one - b1
        bc1 = f32(1.0 - math.pow(float(b1), int(step))) if int(bias_corrs[k]) == 1 else one
        bc2 = f32(1.0 - math.pow(float(b2), int(step))) if int(bias_corrs[k]) == 1 else one
        if decay == 0:
            p = np.zeros_like(p)
        with np.errstate(all="ignore"):
            sg = g / cs
            if mode == 0:
                sg = sg + decay * p
                m = m * b1 + b3 * sg
                v = v * b2 + (one - b2) * sg * sg
                u = (m / bc1) / (np.sqrt(v / bc2) + eps)
            else:
                m = m * b1 + b3 * sg
                v = v * b2 + (one - b2) * sg * sg
                u = (m / bc1) / (np.sqrt(v / bc2) + eps) + decay * p
        _store(lists[4][k], u)
        _store(lists[2][k], m)
        _store(lists[3][k], v)
        ur = _round_trip(u, lists[4][k].dtype).astype(np.float64)
        sums.append(float(np.sum(ur * ur)))
    return np.array(sums, dtype=np.float32)


def update_weights(noop, lists, param_norms, update_norms, offsets, lr, decays, use_nvlamb):
    """multi_tensor_distopt_lamb_kernel.cu:336-389.  lists = [u, p, p_copy]."""
    if noop[0] == 1:
        return
    lr = f32(lr)
    for k in range(len(lists[0])):
        ratio = lr
        if use_nvlamb or f32(decays[k]) != 0:
            pn, un = f32(param_norms[k]), f32(update_norms[int(offsets[k])])
            ratio = lr * (pn / un) if (un != 0 and pn != 0) else lr
        u, p = _to_np(lists[0][k]), _to_np(lists[1][k])
        p = p - f32(ratio) * u
        _store(lists[1][k], p)
        _store(lists[2][k], p)
