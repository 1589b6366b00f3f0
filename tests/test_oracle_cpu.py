"""CPU: pins the oracle (oracle/mt_oracle.py) against the known answers the reference's own tests
hold for this path, and against torch.optim (what the reference's optimizer tests compare to)."""
import math

import numpy as np
import pytest
import torch

from oracle import mt_oracle as O


# ---- tests/L0/run_amp/test_multi_tensor_scale.py:36-77: 4.0 * (1/4) == 1.0, flag semantics ----------
@pytest.mark.parametrize("in_t", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("out_t", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("sizea,sizeb", [(777, 555), (2048 * 32 + 1, 333), (33333, 1)])
def test_scale_kat(in_t, out_t, sizea, sizeb):
    ins = [torch.full((sizea,), 4.0, dtype=in_t), torch.full((sizeb,), 4.0, dtype=in_t)] * 3
    outs = [torch.zeros_like(x, dtype=out_t) for x in ins]
    flag = [0]
    O.multi_tensor_scale(flag, [ins, outs], 1.0 / 4.0)
    assert flag[0] == 0
    for o in outs:
        assert torch.equal(o, torch.ones_like(o))


@pytest.mark.parametrize("val", [float("nan"), float("inf"), -float("inf")])
@pytest.mark.parametrize("t,ind", [(0, 0), (5, -1), (2, 17)])
def test_scale_flag_injection(val, t, ind):
    ins = [torch.full((100,), 4.0) for _ in range(6)]
    outs = [torch.zeros(100) for _ in range(6)]
    ins[t][ind] = val
    flag = [0]
    O.multi_tensor_scale(flag, [ins, outs], 0.25)
    assert flag[0] == 1


# ---- tests/L0/run_amp/test_multi_tensor_axpby.py:34-39: 2*4 + 8*16 == 136 ----------------------------
@pytest.mark.parametrize("x_t,y_t,o_t", [(torch.float32,) * 3, (torch.float16, torch.float32, torch.float16),
                                          (torch.bfloat16, torch.bfloat16, torch.float32)])
def test_axpby_kat(x_t, y_t, o_t):
    xs = [torch.full((555,), 4.0, dtype=x_t), torch.full((33333,), 4.0, dtype=x_t)]
    ys = [torch.full((555,), 16.0, dtype=y_t), torch.full((33333,), 16.0, dtype=y_t)]
    outs = [torch.zeros(555, dtype=o_t), torch.zeros(33333, dtype=o_t)]
    flag = [0]
    O.multi_tensor_axpby(flag, [xs, ys, outs], 2.0, 8.0, -1)
    assert flag[0] == 0
    for o in outs:
        assert torch.equal(o, torch.full_like(o, 136.0))


@pytest.mark.parametrize("arg,which,expect", [(-1, "x", 1), (-1, "y", 1), (0, "x", 1), (0, "y", 0),
                                              (1, "x", 0), (1, "y", 1), (2, "x", 0)])
def test_axpby_arg_to_check(arg, which, expect):
    xs, ys, outs = [torch.ones(64)], [torch.ones(64)], [torch.zeros(64)]
    (xs if which == "x" else ys)[0][3] = float("inf")
    flag = [0]
    O.multi_tensor_axpby(flag, [xs, ys, outs], 1.0, 1.0, arg)
    assert flag[0] == expect


# ---- tests/L0/run_amp/test_multi_tensor_l2norm.py:35-56: constant 4.0 -> 4*sqrt(n) -------------------
@pytest.mark.parametrize("per_tensor", [False, True])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_l2norm_kat(per_tensor, dtype):
    sizea, sizeb, rep = 33333, 555, 3
    xs = [torch.full((sizea,), 4.0, dtype=dtype), torch.full((sizeb,), 4.0, dtype=dtype)] * rep
    flag = [0]
    norm, per = O.multi_tensor_l2norm(2048 * 32, flag, [xs], per_tensor)
    ref = torch.full(((sizea + sizeb) * rep,), 4.0).norm()
    assert math.isclose(float(norm), float(ref), rel_tol=1e-5)
    assert flag[0] == 0
    if per_tensor:
        want = [4.0 * math.sqrt(sizea), 4.0 * math.sqrt(sizeb)] * rep
        np.testing.assert_allclose(per, np.array(want, dtype=np.float32), rtol=1e-5)
    else:
        assert per is None


def test_l2norm_chunk_overflow_flag():
    # finite values whose per-chunk sum of squares overflows fp32 raise the flag
    # (csrc/multi_tensor_l2norm_kernel.cu:103)
    x = [torch.full((1024,), 1e19)]
    flag = [0]
    O.multi_tensor_l2norm(512, flag, [x], False)
    assert flag[0] == 1
    x = [torch.full((1024,), 1e17)]
    flag = [0]
    norm, _ = O.multi_tensor_l2norm(512, flag, [x], False)
    assert flag[0] == 0 and math.isclose(float(norm), 1e17 * 32, rel_tol=1e-5)


def test_l2norm_mp_gated():
    x = [torch.ones(10)]
    flag = [1]
    assert O.multi_tensor_l2norm(64, flag, [x], True, gated=True) == (None, None)


# ---- tests/L0/run_optimizers/test_fused_optimizer.py:63-85: FusedAdam vs torch.optim over 7 steps ----
def _adam_vs_torch(adam_w, wd, amsgrad=False):
    torch.manual_seed(9876)
    shapes = [(4096, 8), (4096,), (1,), (333,)]
    ref_p = [torch.rand(s).requires_grad_() for s in shapes]
    p = [x.detach().clone() for x in ref_p]
    m = [torch.zeros_like(x) for x in p]
    v = [torch.zeros_like(x) for x in p]
    cls = torch.optim.AdamW if adam_w else torch.optim.Adam
    opt = cls(ref_p, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    for step in range(1, 8):
        grads = [torch.rand_like(x) for x in p]
        for rp, g in zip(ref_p, grads):
            rp.grad = g.clone()
        opt.step()
        O.multi_tensor_adam([grads, p, m, v], 5e-4, 0.9, 0.999, 1e-8, step, 1 if adam_w else 0, 1, wd)
        for a, b in zip(p, ref_p):
            assert (a - b.detach()).abs().max() <= 1e-6


def test_adam_l2_mode_vs_torch_adam():
    _adam_vs_torch(False, 0.0)
    _adam_vs_torch(False, 0.1)


def test_adamw_mode_vs_torch_adamw():
    _adam_vs_torch(True, 0.01)


@pytest.mark.parametrize("momentum,nesterov,wd,dampening", [(0.0, False, 0.0, 0.0), (0.9, False, 1e-4, 0.0),
                                                            (0.9, True, 1e-4, 0.0), (0.5, False, 0.0, 0.1)])
def test_sgd_vs_torch(momentum, nesterov, wd, dampening):
    torch.manual_seed(9876)
    ref_p = [torch.rand(5000).requires_grad_(), torch.rand(17).requires_grad_()]
    p = [x.detach().clone() for x in ref_p]
    mom = [torch.zeros_like(x) for x in p]
    opt = torch.optim.SGD(ref_p, lr=0.25, momentum=momentum, nesterov=nesterov, weight_decay=wd, dampening=dampening)
    for it in range(7):
        grads = [torch.rand_like(x) for x in p]
        for rp, g in zip(ref_p, grads):
            rp.grad = g.clone()
        opt.step()
        flag = [0]
        O.multi_tensor_sgd(flag, [[g * 128.0 for g in grads], p, mom], wd, momentum, dampening, 0.25, nesterov,
                           it == 0, False, 1.0 / 128.0)
        for a, b in zip(p, ref_p):
            assert (a - b.detach()).abs().max() <= 1e-5


def test_sgd_noop_skips():
    p, g, mom = [torch.ones(8)], [torch.ones(8)], [torch.zeros(8)]
    O.multi_tensor_sgd([1], [g, p, mom], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0)
    assert torch.equal(p[0], torch.ones(8)) and torch.equal(mom[0], torch.zeros(8))


# ---- LAMB against a direct transcription of the published update (You et al.), one tensor ----------
def test_lamb_single_step_closed_form():
    torch.manual_seed(0)
    n = 1000
    g0, p0 = torch.randn(n), torch.randn(n)
    g, p, m, v = [g0.clone()], [p0.clone()], [torch.zeros(n)], [torch.zeros(n)]
    lr, b1, b2, eps, wd = 1e-2, 0.9, 0.999, 1e-6, 0.01
    gn = float(g0.norm())
    flag = [0]
    O.multi_tensor_lamb(65536, flag, [g, p, m, v], lr, b1, b2, eps, 1, 1, wd, 1, 1, gn, 1.0, False)
    sg = g0.double() / max(gn / 1.0, 1.0)
    m1 = (1 - b1) * sg
    v1 = (1 - b2) * sg * sg
    u = (m1 / (1 - b1)) / ((v1 / (1 - b2)).sqrt() + eps) + wd * p0.double()
    ratio = lr * float(p0.double().norm() / u.norm())
    want = p0.double() - ratio * u
    assert (p[0].double() - want).abs().max() < 1e-5
    assert (g[0].double() - u).abs().max() < 1e-4   # g now holds the update
    assert flag[0] == 0


def test_dist_adam_matches_adam_mode1_when_eps_outside():
    # with per-tensor arrays equal to the scalars, eps_mode 1 and grad_scale 1 the sharded kernel is AdamW
    torch.manual_seed(1)
    n = 777
    g, p = torch.randn(n), torch.randn(n)
    a = [[g.clone()], [p.clone()], [torch.zeros(n)], [torch.zeros(n)]]
    b = [[p.clone()], [torch.zeros(n)], [torch.zeros(n)], [g.clone()]]
    O.multi_tensor_adam(a, 1e-3, 0.9, 0.999, 1e-8, 3, 1, 1, 0.01)
    O.multi_tensor_fused_adam(b, [0.9], [0.999], [1], [1e-8], [0.01], 1e-3, 1.0, 3, 1)
    assert (a[1][0] - b[0][0]).abs().max() < 1e-7


def test_python_fallback_overflow_and_scale():
    g = torch.full((10,), 8.0)
    out = torch.zeros(10)
    assert O.scale_check_overflow_python(g, out, 0.5, True) is False
    assert torch.equal(out, torch.full((10,), 4.0))
    g[3] = float("inf")
    assert O.scale_check_overflow_python(g, out, 0.5, True) is True
    assert O.unscale_python([g], [out], 2.0) is True
