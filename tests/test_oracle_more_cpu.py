"""CPU: oracle restatements of Adagrad / NovoGrad pinned against torch.optim.Adagrad (what the reference's
tests/L0/run_optimizers/test_fused_optimizer.py compares FusedAdagrad with) and a float64 transcription of NovoGrad."""
import numpy as np
import torch

from oracle import mt_oracle as O


def test_adagrad_vs_torch():
    torch.manual_seed(9876)
    ref_p = [torch.rand(5000).requires_grad_(), torch.rand(17).requires_grad_()]
    p = [x.detach().clone() for x in ref_p]
    h = [torch.zeros_like(x) for x in p]
    opt = torch.optim.Adagrad(ref_p, lr=5e-4, eps=1e-8, weight_decay=1e-5)
    for _ in range(7):
        grads = [torch.rand_like(x) for x in p]
        for rp, g in zip(ref_p, grads):
            rp.grad = g.clone()
        opt.step()
        O.multi_tensor_adagrad([grads, p, h], 5e-4, 1e-8, 0, 1e-5)
        for a, b in zip(p, ref_p):
            assert (a - b.detach()).abs().max() <= 1e-6


def test_novograd_single_step_closed_form():
    torch.manual_seed(1)
    g0, p0 = torch.randn(1000), torch.randn(1000)
    g, p, m = [g0.clone()], [p0.clone()], [torch.zeros(1000)]
    norms = np.array([2.0], dtype=np.float32)
    flag = [0]
    lr, b1, b2, eps, wd = 1e-2, 0.95, 0.98, 1e-8, 0.01
    O.multi_tensor_novograd(65536, flag, [g, p, m], norms, lr, b1, b2, eps, 1, 1, wd, 1, 1, 2)
    gn = float(g0.double().norm())
    v = (b2 * 2.0 ** 2 + (1 - b2) * gn ** 2) ** 0.5
    assert abs(float(norms[0]) - v) <= 1e-5 * v
    m1 = (1 - b1) * g0.double()
    want = p0.double() - lr * ((m1 / (1 - b1)) / (v / (1 - b2) ** 0.5 + eps) + wd * p0.double())
    assert (p[0].double() - want).abs().max() < 1e-6 and flag[0] == 0


def test_norm_out_linf_drops_nan_and_flags_inf():
    out = np.array([1.0, 1.0], dtype=np.float32)
    x = [torch.tensor([1.0, float("nan"), -3.0]), torch.tensor([2.0, float("inf")])]
    flag = [0]
    O.multi_tensor_norm_out(64, flag, [x], out, 0.5, 0.5, 0)
    assert out[0] == 0.5 * 1.0 + 0.5 * 3.0 and flag[0] == 1 and np.isinf(out[1])
