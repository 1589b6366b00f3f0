"""GPU: fused_adam_cuda.reversible_adam / maybe_adam_undo (C ABI b200mt_reversible_adam / b200mt_maybe_adam_undo) against
the reference's own kernels (oracle/_ref/fused_adam_cuda_ref.so, apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:570-732)
and the CPU oracle.  The undo test also pins the one conscious divergence: the reference kernel reverts elements 0 mod 4
only; here every element is reverted."""
import numpy as np
import pytest
import torch

from oracle import mt_oracle as O
from util import assert_close_rel, load_ref

pytestmark = pytest.mark.gpu
HYPER = dict(lr=1e-2, beta1=0.9, beta2=0.99, eps=1e-8, grad_scale=8.0, step=3, mode=1, bias_correction=1, decay=0.01)


def flag_tensor(v=0):
    return torch.full((1,), v, dtype=torch.int32, device="cuda")


@pytest.fixture(scope="module")
def refm():
    m = load_ref("fused_adam_cuda_ref")
    if m is None:
        pytest.skip("oracle/_ref/fused_adam_cuda_ref.so not built")
    return m


def make(n, gdt, cdt, seed, bad=()):
    g = torch.Generator(device="cuda").manual_seed(seed)
    p = torch.randn(n, generator=g, device="cuda") * 0.1
    m = torch.randn(n, generator=g, device="cuda") * 0.01
    v = torch.rand(n, generator=g, device="cuda") * 1e-3
    gr = (torch.randn(n, generator=g, device="cuda") * 8.0).to(gdt)
    for i in bad:
        gr[i] = float("inf") if i % 2 == 0 else float("nan")
    pc = torch.zeros(n if cdt is not None else 0, dtype=cdt or gdt, device="cuda")
    return p, pc, m, v, gr


def args(h):
    return (h["lr"], h["beta1"], h["beta2"], h["eps"], h["grad_scale"], h["step"], h["mode"], h["bias_correction"], h["decay"])


@pytest.mark.parametrize("gdt,cdt", [(torch.float16, torch.float16), (torch.bfloat16, torch.bfloat16), (torch.float16, torch.uint8),
                                     (torch.float16, None), (torch.float32, None)])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("bad", [(), (5,), (0, 777)])
def test_reversible_adam_matches_ref_and_oracle(refm, gdt, cdt, mode, bad):
    import fused_adam_cuda as ours
    h = dict(HYPER, mode=mode)
    for n in (1, 7, 555, 65537 + 3):
        b = tuple(i for i in bad if i < n)
        a = make(n, gdt, cdt, 80 + n % 7, b)
        r = tuple(t.clone() for t in a)
        c = tuple(t.cpu().clone() for t in a)
        ours.reversible_adam(*a, *args(h))
        refm.reversible_adam(*r, *args(h))
        ov = O.reversible_adam(c[0], c[1] if cdt is not None else None, c[2], c[3], c[4], *args(h))
        assert ov == bool(b)
        for k, name in ((0, "p"), (2, "m"), (3, "v")):
            assert_close_rel(a[k], r[k], rtol=2e-6, scale="self", what="%s vs reference kernel" % name)
            assert_close_rel(a[k], c[k], rtol=2e-6, scale="self", what="%s vs oracle" % name)
        for i in b:     # a non-finite gradient leaves its element alone
            o = make(n, gdt, cdt, 80 + n % 7, b)
            assert a[0][i] == o[0][i] and a[2][i] == o[2][i] and a[3][i] == o[3][i]
        if cdt is None:
            continue
        pc = a[1]
        if b:           # the marker: +inf in the first element, whatever element overflowed
            assert bool(torch.isinf(pc[:1].float())) if cdt != torch.uint8 else int(pc[0]) == 0x7C
            assert bool(torch.isinf(r[1][:1].float())) if cdt != torch.uint8 else int(r[1][0]) == 0x7C
            f = flag_tensor(0)
            ours.strided_check_finite(f, pc, 1, 1)
            assert int(f) == 1
        lo = 1 if b else 0
        if cdt == torch.uint8:   # bytes: the reference-pinned packing of this kernel's own fp32 result
            assert np.array_equal(pc[lo:].cpu().numpy(), O.pack_e5m2(a[0][lo:].cpu().numpy()))
        else:
            assert torch.equal(pc[lo:], a[0][lo:].to(cdt))


@pytest.mark.parametrize("gdt", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("mode", [0, 1])
def test_undo_restores_every_element_and_matches_the_reference_where_it_reaches(refm, gdt, mode):
    import fused_adam_cuda as ours
    h = dict(HYPER, mode=mode)
    n = 65537 + 3
    bad = (4, 1001)
    p0, pc, m0, v0, g = make(n, gdt, None, 91, bad)
    p, m, v = p0.clone(), m0.clone(), v0.clone()
    ours.reversible_adam(p, pc, m, v, g, *args(h))
    stepped = (p.clone(), m.clone(), v.clone())
    assert not torch.equal(p, p0)
    # flag == 0: nothing moves
    ours.maybe_adam_undo(flag_tensor(0), p, m, v, g, *args(h))
    assert torch.equal(p, stepped[0]) and torch.equal(m, stepped[1]) and torch.equal(v, stepped[2])
    # the reference kernel and both oracle variants on the same stepped state
    rp, rm, rv = (t.clone() for t in stepped)
    refm.maybe_adam_undo(flag_tensor(1), rp, rm, rv, g, *args(h))
    c4 = [t.cpu().clone() for t in stepped]
    call = [t.cpu().clone() for t in stepped]
    O.maybe_adam_undo([1], c4[0], c4[1], c4[2], g.cpu(), *args(h), only_every_fourth=True)
    O.maybe_adam_undo([1], call[0], call[1], call[2], g.cpu(), *args(h))
    ours.maybe_adam_undo(flag_tensor(1), p, m, v, g, *args(h))
    idx = torch.arange(n, device="cuda")
    reached = (idx % 4 == 0)
    for mine, ref, c_ref, c_all, st, name in ((p, rp, c4[0], call[0], stepped[0], "p"), (m, rm, c4[1], call[1], stepped[1], "m"),
                                              (v, rv, c4[2], call[2], stepped[2], "v")):
        # undoing subtracts terms as large as the stepped value: that, not the (smaller) result, sets the rounding error
        sc = float(st.abs().max())
        # the oracle restatement of the kernel AS WRITTEN reproduces the reference on every element ...
        assert_close_rel(ref, c_ref, rtol=2e-6, atol=1e-9, scale=sc, what="reference undo vs oracle(as written) " + name)
        # ... which leaves three elements out of four at their stepped values
        assert torch.equal(ref[~reached], st[~reached])
        # ours agrees with the reference where the reference reaches, and with the full-inverse oracle everywhere
        assert_close_rel(mine[reached], ref[reached], rtol=2e-6, atol=1e-9, scale=sc, what="undo vs reference (0 mod 4) " + name)
        assert_close_rel(mine, c_all, rtol=2e-6, atol=1e-9, scale=sc, what="undo vs oracle " + name)
    # round trip: the state before the step is back (v loses bits to cancellation: absolute bound)
    assert float((p - p0).abs().max()) <= 2e-6
    assert float((m - m0).abs().max()) <= 2e-6
    assert float((v - v0).abs().max()) <= 2e-5
    for i in bad:
        assert p[i] == p0[i] and m[i] == m0[i] and v[i] == v0[i]
    # an empty flag tensor means "always undo" (overflow_flag.numel() ? ptr : NULL, :1003)
    p2, m2, v2 = (t.clone() for t in stepped)
    ours.maybe_adam_undo(torch.zeros(0, dtype=torch.int32, device="cuda"), p2, m2, v2, g, *args(h))
    assert torch.equal(p2, p) and torch.equal(m2, m) and torch.equal(v2, v)


def test_argument_checks_mirror_the_reference():
    import fused_adam_cuda as ours
    p, pc, m, v, g = make(16, torch.float16, torch.float16, 3)
    with pytest.raises(RuntimeError, match="number of elements in m and p"):
        ours.reversible_adam(p, pc, m[:8], v, g, *args(HYPER))
    with pytest.raises(RuntimeError, match="p_copy"):
        ours.reversible_adam(p, pc[:8], m, v, g, *args(HYPER))
    with pytest.raises(RuntimeError, match="float type"):
        ours.reversible_adam(p.half(), pc, m.half(), v.half(), g, *args(HYPER))
    with pytest.raises(RuntimeError, match="number of elements in g and p"):
        ours.maybe_adam_undo(flag_tensor(1), p, m, v, g[:8], *args(HYPER))
