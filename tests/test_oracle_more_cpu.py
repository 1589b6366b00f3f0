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


def test_e5m2_pack_unpack_properties():
    """fused_adam_cuda_kernel.cu:395-463: a byte is the high half of a Half; packing rounds half up in magnitude.
    (Pinned bit-for-bit against the reference kernel in tests/test_parity_more_gpu.py.)"""
    b = np.arange(256, dtype=np.uint8)
    f = O.unpack_e5m2(b)
    assert f[0x3C] == 1.0 and f[0x3D] == 1.25 and f[0xBC] == -1.0 and np.isinf(f[0x7C]) and np.isnan(f[0x7E])
    fin = np.isfinite(f)
    assert np.array_equal(O.pack_e5m2(f[fin]), b[fin])                    # representable values survive
    x = np.array([1.0, 1.124, 1.125, 1.126, 1.374, 1.375, -1.125, 0.0, 65504.0, 61440.0, 1e-9, np.inf, np.nan], dtype=np.float32)
    want = np.array([0x3C, 0x3C, 0x3D, 0x3D, 0x3D, 0x3E, 0xBD, 0x00, 0x7C, 0x7C, 0x00, 0x7C, 0x7F], dtype=np.uint8)
    assert np.array_equal(O.pack_e5m2(x), want)
    t = torch.zeros(4, dtype=torch.uint8)
    O.maybe_cast_mt(None, [[torch.tensor([1.0, -2.5, 0.3, 100.0])], [t]])
    out = torch.zeros(4)
    O.maybe_cast_mt(None, [[t], [out]])
    assert torch.equal(out, torch.tensor([1.0, -2.5, 0.3125, 96.0]))


def test_reversible_adam_and_its_undo():
    """fused_adam_cuda_kernel.cu:570-732: non-finite gradients are skipped and flagged through p_copy[0]; the undo is the
    inverse of the step -- on every element, or (only_every_fourth) on the elements the kernel as written reaches."""
    g = torch.Generator().manual_seed(5)
    n = 1000
    p0, m0, v0 = torch.randn(n, generator=g) * 0.1, torch.randn(n, generator=g) * 0.01, torch.rand(n, generator=g) * 1e-3
    gr = (torch.randn(n, generator=g) * 8).half()
    gr[7] = float("inf")
    h = (1e-2, 0.9, 0.99, 1e-8, 8.0, 3, 1, 1, 0.01)
    p, m, v, pc = p0.clone(), m0.clone(), v0.clone(), torch.zeros(n, dtype=torch.float16)
    assert O.reversible_adam(p, pc, m, v, gr, *h) is True
    assert bool(torch.isinf(pc[0])) and torch.equal(pc[1:], p[1:].half())
    assert p[7] == p0[7] and m[7] == m0[7] and v[7] == v0[7]
    # element 3 against the formula in double
    ss = 1e-2 * (1 - 0.99 ** 3) ** 0.5 / (1 - 0.9 ** 3)
    sg = float(gr[3]) / 8.0
    m3 = 0.9 * float(m0[3]) + 0.1 * sg
    v3 = 0.99 * float(v0[3]) + 0.01 * sg * sg
    p3 = float(p0[3]) - ss * (m3 / (v3 ** 0.5 + 1e-8) + 0.01 * float(p0[3]))
    assert abs(float(p[3]) - p3) <= 1e-6 and abs(float(m[3]) - m3) <= 1e-7 and abs(float(v[3]) - v3) <= 1e-7
    stepped = (p.clone(), m.clone(), v.clone())
    O.maybe_adam_undo([0], p, m, v, gr, *h)                                # flag clear: nothing happens
    assert torch.equal(p, stepped[0])
    q = [t.clone() for t in stepped]
    O.maybe_adam_undo([1], q[0], q[1], q[2], gr, *h, only_every_fourth=True)
    idx = torch.arange(n)
    assert torch.equal(q[0][idx % 4 != 0], stepped[0][idx % 4 != 0])
    O.maybe_adam_undo([1], p, m, v, gr, *h)
    assert torch.equal(q[0][idx % 4 == 0], p[idx % 4 == 0])
    assert float((p - p0).abs().max()) <= 2e-6 and float((m - m0).abs().max()) <= 2e-6 and float((v - v0).abs().max()) <= 2e-5
