"""GPU parity: the sm_100a kernels, called through the drop-in modules (amp_C / distributed_adam_cuda /
fused_adam_cuda -> C ABI of libb200mt.so), against
  (1) the CPU oracle (oracle/mt_oracle.py) and
  (2) the reference's own CUDA kernels compiled unmodified for sm_100a (oracle/_ref/*.so),
on identical seeded inputs.  The same runs also pin the oracle against the reference.

Tolerances (BASELINE.json north_star): overflow flag / noop skipping bit-exact; fp32 Adam/SGD
1e-6 relative; L2 norms and LAMB trust ratios 1e-5 relative.  Second moments (v: a sum of
non-negative terms) and norms are held to PURE element-wise relative error, |ours - want| <=
rtol*|want| (assert_rel_pure / assert_close_rel without `scale`).  For p, m and stored updates the
bound is |ours - want| <= rtol*|want| + rtol*max|want| over the tensor: the second term covers elements
that come out of a cancellation (m = b1*m + (1-b1)*g landing near zero), where two correct fp32
evaluations differ by an ulp of the operands (FMA contraction; the reference is built with
--use_fast_math: approximate division / sqrt, denormals flushed), not of the tiny result.  The
applied Adam update (p_new - p_old) is checked separately.  Scale / cast results are bit-exact.
tests/test_parity_fullsize_gpu.py repeats the comparison at the BASELINE sizes and records the error distribution.
"""
import itertools
import math

import numpy as np
import pytest
import torch

from oracle import mt_oracle as O
from util import assert_close_rel, assert_rel_pure, load_ref

pytestmark = pytest.mark.gpu

CHUNK = 2048 * 32
DT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


@pytest.fixture(scope="module")
def amp_C():
    import amp_C as m
    return m


@pytest.fixture(scope="module")
def ref():
    m = load_ref("amp_C_ref")
    if m is None:
        pytest.skip("oracle/_ref/amp_C_ref.so not built")
    return m


def flag_tensor(v=0):
    return torch.full((1,), v, dtype=torch.int32, device="cuda")


def cpu(lists):
    return [[t.detach().cpu().clone() for t in l] for l in lists]


def clone(lists):
    return [[t.detach().clone() for t in l] for l in lists]


def rand_lists(sizes, dtypes, seed, scale=1.0, positive=()):
    g = torch.Generator(device="cuda").manual_seed(seed)
    out = []
    for d, dt in enumerate(dtypes):
        lst = []
        for n in sizes:
            x = torch.randn(n, generator=g, device="cuda", dtype=torch.float32) * scale
            if d in positive:
                x = x.abs()
            lst.append(x.to(dt))
        out.append(lst)
    return out


SIZES = [1, 3, 555, 2048 * 32 + 1, 33333, 8192, 16, 2048 * 32 * 2 - 1]


# =================================================================================================
# scale
# =================================================================================================
@pytest.mark.parametrize("in_t,out_t", list(itertools.product(DT, DT)))
@pytest.mark.parametrize("chunk", [CHUNK, 333])
def test_scale_matches_oracle_and_ref(amp_C, ref, in_t, out_t, chunk):
    ins, outs = rand_lists(SIZES, [DT[in_t], DT[out_t]], seed=1, scale=100.0)
    c_in, c_out = cpu([ins, outs])
    r_in, r_out = clone([ins, outs])
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    amp_C.multi_tensor_scale(chunk, f, [ins, outs], 1.0 / 128.0)
    ref.multi_tensor_scale(chunk, rf, [r_in, r_out], 1.0 / 128.0)
    O.multi_tensor_scale(of, [c_in, c_out], 1.0 / 128.0)
    assert int(f) == int(rf) == of[0] == 0
    for a, b, c in zip(outs, r_out, c_out):
        assert torch.equal(a, b)            # one multiply + one rounding: bit-exact
        assert torch.equal(a.cpu(), c)


@pytest.mark.parametrize("val", [float("nan"), float("inf"), -float("inf")])
@pytest.mark.parametrize("t,ind", [(0, 0), (-1, -1), (3, 40000), (2, 554)])
@pytest.mark.parametrize("in_t", ["f32", "f16"])
def test_scale_overflow_flag(amp_C, ref, val, t, ind, in_t):
    ins, outs = rand_lists(SIZES, [DT[in_t], torch.float32], seed=2)
    ins[t][ind] = val
    r_in, r_out = clone([ins, outs])
    f, rf = flag_tensor(), flag_tensor()
    amp_C.multi_tensor_scale(CHUNK, f, [ins, outs], 0.5)
    ref.multi_tensor_scale(CHUNK, rf, [r_in, r_out], 0.5)
    assert int(f) == 1 and int(rf) == 1
    for a, b in zip(outs, r_out):       # values still propagate (the kernel does not early-exit)
        assert torch.equal(a.isnan(), b.isnan()) and torch.equal(a.nan_to_num(), b.nan_to_num())


def test_scale_inplace_and_reference_kat(amp_C):
    # tests/L0/run_amp/test_multi_tensor_scale.py:36-56: 4.0 * (1/4) == 1.0, flag 0, in place
    xs = [torch.full((n,), 4.0, device="cuda") for _ in range(55) for n in (777, 2048 * 32 + 1, 33333)]
    f = flag_tensor()
    amp_C.multi_tensor_scale(CHUNK, f, [xs, xs], 0.25)
    assert int(f) == 0
    assert all(torch.equal(x, torch.ones_like(x)) for x in xs)


def test_scale_misaligned_views_and_empty(amp_C, ref):
    base_i = torch.randn(100003, device="cuda")
    base_o = torch.zeros(100003, device="cuda", dtype=torch.float16)
    ins = [base_i[1:50001], base_i[50002:50002], base_i[50003:100003], torch.randn(7, device="cuda")[3:]]
    outs = [base_o[3:50003], base_o[0:0], base_o[50003:100003], torch.zeros(7, device="cuda", dtype=torch.float16)[1:5]]
    r_out = [torch.zeros_like(o) for o in outs]
    f, rf = flag_tensor(), flag_tensor()
    amp_C.multi_tensor_scale(CHUNK, f, [ins, outs], 3.0)
    ref.multi_tensor_scale(CHUNK, rf, [ins, r_out], 3.0)
    for a, b in zip(outs, r_out):
        assert torch.equal(a, b)
    assert base_o[0:3].abs().sum() == 0 and base_o[50003 - 0:50003].numel() == 0


def test_scale_many_tensors_single_launch(amp_C):
    import _b200mt
    big = torch.randn(10000 * 1024, device="cuda")
    ins = list(big.view(10000, 1024).unbind(0))
    out_big = torch.empty_like(big)
    outs = list(out_big.view(10000, 1024).unbind(0))
    f = flag_tensor()
    before = _b200mt.lib.b200mt_launch_count()
    amp_C.multi_tensor_scale(CHUNK, f, [ins, outs], 2.0)
    assert _b200mt.lib.b200mt_launch_count() - before == 1   # the reference needs 157 launches here
    assert torch.equal(out_big, big * 2.0) and int(f) == 0


def test_error_messages_match_reference(amp_C):
    f = flag_tensor()
    a, b = torch.ones(4, device="cuda"), torch.ones(5, device="cuda")
    with pytest.raises(RuntimeError, match="Size mismatch"):
        amp_C.multi_tensor_scale(CHUNK, f, [[a], [b]], 1.0)
    with pytest.raises(RuntimeError, match="Size mismatch among tensor lists"):
        amp_C.multi_tensor_scale(CHUNK, f, [[a, a], [a]], 1.0)
    with pytest.raises(RuntimeError, match="is not > 0"):
        amp_C.multi_tensor_scale(CHUNK, f, [[], []], 1.0)
    with pytest.raises(RuntimeError, match="expected input to be on cuda"):
        amp_C.multi_tensor_scale(CHUNK, f, [[a.cpu()], [a.cpu()]], 1.0)
    with pytest.raises(RuntimeError, match="not contiguous"):
        nc = torch.ones(8, 8, device="cuda").t()
        amp_C.multi_tensor_scale(CHUNK, f, [[nc], [nc]], 1.0)
    with pytest.raises(RuntimeError, match="depth"):
        amp_C.multi_tensor_scale(CHUNK, f, [[a]], 1.0)
    with pytest.raises(RuntimeError, match="not implemented for"):
        i = torch.ones(4, device="cuda", dtype=torch.float64)
        amp_C.multi_tensor_scale(CHUNK, f, [[i], [i]], 1.0)


# =================================================================================================
# axpby
# =================================================================================================
@pytest.mark.parametrize("x_t,y_t,o_t", list(itertools.product(DT, DT, DT)))
def test_axpby_matches_oracle_and_ref(amp_C, ref, x_t, y_t, o_t):
    xs, ys, outs = rand_lists(SIZES, [DT[x_t], DT[y_t], DT[o_t]], seed=3, scale=10.0)
    c = cpu([xs, ys, outs])
    r = clone([xs, ys, outs])
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    amp_C.multi_tensor_axpby(CHUNK, f, [xs, ys, outs], 2.0, 0.37, -1)
    ref.multi_tensor_axpby(CHUNK, rf, r, 2.0, 0.37, -1)
    O.multi_tensor_axpby(of, c, 2.0, 0.37, -1)
    assert int(f) == int(rf) == of[0] == 0
    ulp = {"f32": 2e-7, "f16": 1e-3, "bf16": 8e-3}[o_t]   # FMA contraction may differ by one rounding
    for a, b, cc in zip(outs, r[2], c[2]):
        assert_close_rel(a, b, rtol=ulp, scale="self", what="vs ref")
        assert_close_rel(a, cc, rtol=ulp, scale="self", what="vs oracle")


def test_axpby_reference_kat_and_checks(amp_C, ref):
    # tests/L0/run_amp/test_multi_tensor_axpby.py:34-39: 2*4 + 8*16 == 136
    xs = [torch.full((n,), 4.0, device="cuda") for n in (555, 33333)]
    ys = [torch.full((n,), 16.0, device="cuda", dtype=torch.float16) for n in (555, 33333)]
    outs = [torch.zeros(n, device="cuda") for n in (555, 33333)]
    f = flag_tensor()
    amp_C.multi_tensor_axpby(CHUNK, f, [xs, ys, outs], 2.0, 8.0, -1)
    assert int(f) == 0 and all(torch.equal(o, torch.full_like(o, 136.0)) for o in outs)
    for arg, which, expect in [(-1, 0, 1), (-1, 1, 1), (0, 0, 1), (0, 1, 0), (1, 0, 0), (1, 1, 1), (2, 0, 0)]:
        lists = [[torch.ones(4099, device="cuda")], [torch.ones(4099, device="cuda")], [torch.zeros(4099, device="cuda")]]
        lists[which][0][4098] = float("nan")
        f, rf = flag_tensor(), flag_tensor()
        amp_C.multi_tensor_axpby(CHUNK, f, lists, 1.0, 1.0, arg)
        ref.multi_tensor_axpby(CHUNK, rf, clone(lists), 1.0, 1.0, arg)
        assert int(f) == int(rf) == expect, (arg, which)


# =================================================================================================
# l2norm family
# =================================================================================================
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
@pytest.mark.parametrize("per_tensor", [False, True])
@pytest.mark.parametrize("chunk", [CHUNK, 333, 33333])
def test_l2norm_matches_oracle_and_ref(amp_C, ref, dt, per_tensor, chunk):
    (xs,) = rand_lists(SIZES + [300001], [DT[dt]], seed=4, scale=3.0)
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    n, p = amp_C.multi_tensor_l2norm(chunk, f, [xs], per_tensor)
    rn, rp = ref.multi_tensor_l2norm(chunk, rf, [xs], per_tensor)
    on, op = O.multi_tensor_l2norm(chunk, of, cpu([xs]), per_tensor)
    assert int(f) == int(rf) == of[0] == 0
    assert n.shape == rn.shape == (1,) and p.shape == rp.shape
    assert_close_rel(n, rn, rtol=1e-5, what="norm vs ref")
    assert math.isclose(float(n), float(on), rel_tol=1e-5)
    if per_tensor:
        assert_close_rel(p, rp, rtol=1e-5, what="per-tensor vs ref")
        assert_close_rel(p, torch.from_numpy(op), rtol=1e-5, what="per-tensor vs oracle")


def test_l2norm_reference_kat(amp_C):
    # tests/L0/run_amp/test_multi_tensor_l2norm.py:35-56
    for sizea, sizeb, rep in [(7777 * 77, 555, 1), (2048 * 32 + 1, 333, 55)]:
        xs = [torch.full((sizea,), 4.0, device="cuda"), torch.full((sizeb,), 4.0, device="cuda")] * rep
        f = flag_tensor()
        n, p = amp_C.multi_tensor_l2norm(CHUNK, f, [xs], True)
        want = torch.full(((sizea + sizeb) * rep,), 4.0, device="cuda").norm()
        assert torch.allclose(n, want.reshape(1)) and int(f) == 0
        assert torch.allclose(p, torch.tensor([4.0 * math.sqrt(sizea), 4.0 * math.sqrt(sizeb)] * rep, device="cuda"))


def test_l2norm_is_bitwise_deterministic(amp_C):
    (xs,) = rand_lists([1 << 20, 12345, 7] * 5, [torch.float32], seed=5)
    outs = []
    for _ in range(4):
        n, p = amp_C.multi_tensor_l2norm(CHUNK, flag_tensor(), [xs], True)
        outs.append((n.clone(), p.clone()))
    for n, p in outs[1:]:
        assert torch.equal(n, outs[0][0]) and torch.equal(p, outs[0][1])


@pytest.mark.parametrize("val,expect", [(float("inf"), 1), (float("nan"), 1), (1e19, 1), (1e17, 0)])
def test_l2norm_flag_semantics(amp_C, ref, val, expect):
    # chunk-sum overflow of finite values raises the flag too (csrc/multi_tensor_l2norm_kernel.cu:103)
    xs = [torch.ones(70000, device="cuda"), torch.full((1024,), 1.0, device="cuda")]
    xs[1][:] = val if val in (1e19, 1e17) else 1.0
    if val not in (1e19, 1e17):
        xs[0][69999] = val
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    amp_C.multi_tensor_l2norm(512, f, [xs], True)
    ref.multi_tensor_l2norm(512, rf, [xs], True)
    O.multi_tensor_l2norm(512, of, cpu([xs]), True)
    assert int(f) == int(rf) == of[0] == expect


def test_l2norm_mp_gating(amp_C, ref):
    (xs,) = rand_lists(SIZES, [torch.float16], seed=6)
    f = flag_tensor(0)
    n, p = amp_C.multi_tensor_l2norm_mp(CHUNK, f, [xs], True)
    rn, rp = ref.multi_tensor_l2norm_mp(CHUNK, flag_tensor(0), [xs], True)
    assert_close_rel(n, rn, rtol=1e-5)
    assert_close_rel(p, rp, rtol=1e-5)
    # flag already set: nothing runs, nothing is written, later calls still work (ticket stays consistent)
    f1 = flag_tensor(1)
    amp_C.multi_tensor_l2norm_mp(CHUNK, f1, [xs], True)
    assert int(f1) == 1
    n2, _ = amp_C.multi_tensor_l2norm_mp(CHUNK, flag_tensor(0), [xs], True)
    assert torch.equal(n2, n)


@pytest.mark.parametrize("in_t,out_t", [("f32", "f32"), ("f16", "f32"), ("f32", "f16")])
def test_l2norm_scale(amp_C, ref, in_t, out_t):
    ins, outs = rand_lists(SIZES, [DT[in_t], DT[out_t]], seed=7)
    r_out = [torch.zeros_like(o) for o in outs]
    f, rf = flag_tensor(), flag_tensor()
    n, p = amp_C.multi_tensor_l2norm_scale(CHUNK, f, [ins, outs], 0.125, True)
    rn, rp = ref.multi_tensor_l2norm_scale(CHUNK, rf, [ins, r_out], 0.125, True)
    assert_close_rel(n, rn, rtol=1e-5)
    assert_close_rel(p, rp, rtol=1e-5)
    for a, b in zip(outs, r_out):
        assert torch.equal(a, b)
    assert int(f) == int(rf) == 0


# =================================================================================================
# adam
# =================================================================================================
def _adam_state(dt, seed, sizes=SIZES):
    g, p, m, v = rand_lists(sizes, [dt] * 4, seed=seed, positive=(3,))
    for x in m:
        x.mul_(0.1)
    for x in v:
        x.mul_(0.01)
    return [g, p, m, v]


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("bias_correction", [0, 1])
@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_adam_fp32_matches_ref_and_oracle(amp_C, ref, mode, bias_correction, wd):
    lists = _adam_state(torch.float32, seed=10)
    r, c = clone(lists), cpu(lists)
    f = flag_tensor()
    for step in (1, 2, 7):
        args = (1e-3, 0.9, 0.999, 1e-8, step, mode, bias_correction, wd)
        p_old = [x.clone() for x in lists[1]]
        rp_old = [x.clone() for x in r[1]]
        amp_C.multi_tensor_adam(CHUNK, f, lists, *args)
        ref.multi_tensor_adam(CHUNK, f, r, *args)
        O.multi_tensor_adam(c, *args)
        # the applied update itself (p_new - p_old), as far as fp32 storage of p lets it be observed
        for a, a0, b, b0 in zip(lists[1], p_old, r[1], rp_old):
            assert_close_rel(a - a0, b - b0, rtol=2e-4, scale="self", what="applied update vs ref step %d" % step)
        # north_star: fp32 Adam within 1e-6 relative (absolute floor for the fast-math reference)
        for k, name in ((1, "p"), (2, "m")):
            for a, b, cc in zip(lists[k], r[k], c[k]):
                assert_close_rel(a, b, rtol=1e-6, scale="self", what="%s vs ref step %d" % (name, step))
                assert_close_rel(a, cc, rtol=1e-6, scale="self", what="%s vs oracle step %d" % (name, step))
        assert_rel_pure(lists[3], r[3], 1e-6, "v vs ref step %d" % step)                     # no floor for v
        assert_rel_pure(lists[3], [x.cuda() for x in c[3]], 1e-6, "v vs oracle step %d" % step)
    assert int(f) == 0
    for a, b in zip(lists[0], r[0]):
        assert torch.equal(a, b)    # g is never written


@pytest.mark.parametrize("dt", ["f16", "bf16"])
def test_adam_16bit_single_dtype(amp_C, ref, dt):
    lists = _adam_state(DT[dt], seed=11)
    r, c = clone(lists), cpu(lists)
    args = (1e-2, 0.9, 0.999, 1e-6, 3, 1, 1, 0.01)
    amp_C.multi_tensor_adam(CHUNK, flag_tensor(), lists, *args)
    ref.multi_tensor_adam(CHUNK, flag_tensor(), r, *args)
    O.multi_tensor_adam(c, *args)
    ulp = {"f16": 1e-3, "bf16": 8e-3}[dt]   # results are rounded to the storage type: allow one ulp
    for k in (1, 2, 3):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=ulp, scale="self", what="vs ref")
            assert_close_rel(a, cc, rtol=ulp, scale="self", what="vs oracle")


def test_adam_fp64_lists(amp_C, ref):
    lists = _adam_state(torch.float64, seed=12, sizes=[5, 4099, 70001])
    r = clone(lists)
    args = (1e-3, 0.9, 0.999, 1e-8, 2, 1, 1, 0.0)
    amp_C.multi_tensor_adam(CHUNK, flag_tensor(), lists, *args)
    ref.multi_tensor_adam(CHUNK, flag_tensor(), r, *args)
    for k in (1, 2, 3):
        for a, b in zip(lists[k], r[k]):
            assert_close_rel(a, b, rtol=1e-6, scale="self")


@pytest.mark.parametrize("gdt,cdt", [("f16", "f16"), ("bf16", "bf16"), ("f32", "bf16")])
def test_adam_mixed_precision_single_pass(amp_C, gdt, cdt):
    # extension: [g16, p32, m32, v32, p_copy16] in one pass == fp32 Adam on the widened grads + a cast
    sizes = SIZES
    (g,) = rand_lists(sizes, [DT[gdt]], seed=13)
    _, p, m, v = _adam_state(torch.float32, seed=14)
    pc = [torch.zeros(n, device="cuda", dtype=DT[cdt]) for n in sizes]
    base = [[x.float() for x in g], [x.clone() for x in p], [x.clone() for x in m], [x.clone() for x in v]]
    args = (1e-3, 0.9, 0.999, 1e-8, 5, 1, 1, 0.01)
    amp_C.multi_tensor_adam(CHUNK, flag_tensor(), [g, p, m, v, pc], *args)
    amp_C.multi_tensor_adam(CHUNK, flag_tensor(), base, *args)
    for k in (1, 2, 3):
        for a, b in zip([g, p, m, v][k], base[k]):
            assert torch.equal(a, b)
    for a, b in zip(pc, p):
        assert torch.equal(a, b.to(DT[cdt]))


# =================================================================================================
# sgd
# =================================================================================================
SGD_CASES = [("f16", "f16", 3), ("f32", "f32", 3), ("f16", "f32", 4), ("f32", "f32", 4), ("bf16", "bf16", 3), ("bf16", "f32", 4)]


@pytest.mark.parametrize("gdt,wdt,n", SGD_CASES)
@pytest.mark.parametrize("momentum,nesterov,wd,wd_after,dampening", [(0.0, False, 0.0, False, 0.0), (0.9, False, 1e-4, False, 0.0),
                                                                      (0.9, True, 1e-4, True, 0.0), (0.5, False, 0.0, False, 0.1)])
def test_sgd_matches_ref_and_oracle(amp_C, ref, gdt, wdt, n, momentum, nesterov, wd, wd_after, dampening):
    g, = rand_lists(SIZES, [DT[gdt]], seed=20, scale=64.0)
    p, mom = rand_lists(SIZES, [DT[wdt]] * 2, seed=21)
    lists = [g, p, mom]
    if n == 4:
        lists.append([torch.zeros(s, device="cuda", dtype=torch.float16) for s in SIZES])
    r, c = clone(lists), cpu(lists)
    for first_run in (True, False):
        args = (wd, momentum, dampening, 0.1, nesterov, first_run, wd_after, 1.0 / 64.0)
        f, rf, of = flag_tensor(), flag_tensor(), [0]
        amp_C.multi_tensor_sgd(CHUNK, f, lists, *args)
        ref.multi_tensor_sgd(CHUNK, rf, r, *args)
        O.multi_tensor_sgd(of, c, *args)
        rt = 1e-6 if wdt == "f32" else {"f16": 1e-3, "bf16": 8e-3}[wdt]
        for k in range(1, n):
            for a, b, cc in zip(lists[k], r[k], c[k]):
                kt = 1e-3 if k == 3 else rt
                assert_close_rel(a, b, rtol=kt, scale="self", what="list %d vs ref" % k)
                assert_close_rel(a, cc, rtol=kt, scale="self", what="list %d vs oracle" % k)


def test_sgd_noop_flag_skips_bit_exactly(amp_C, ref):
    lists = rand_lists(SIZES, [torch.float16, torch.float32, torch.float32], seed=22) + \
        [[torch.zeros(s, device="cuda", dtype=torch.float16) for s in SIZES]]
    before = clone(lists)
    amp_C.multi_tensor_sgd(CHUNK, flag_tensor(1), lists, 1e-4, 0.9, 0.0, 0.1, False, False, False, 1.0)
    for l0, l1 in zip(before, lists):
        for a, b in zip(l0, l1):
            assert torch.equal(a, b)


def test_sgd_errors(amp_C):
    g, p, mom = rand_lists([16], [torch.float16, torch.float32, torch.float32], seed=23)
    with pytest.raises(RuntimeError, match="Additional output tensors should always be fp16"):
        amp_C.multi_tensor_sgd(CHUNK, flag_tensor(), [g, p, mom, [torch.zeros(16, device="cuda")]], 0, 0, 0, 0.1, False, False, False, 1.0)
    with pytest.raises(RuntimeError, match="only supports some combinations"):
        amp_C.multi_tensor_sgd(CHUNK, flag_tensor(), [g, p, mom], 0, 0, 0, 0.1, False, False, False, 1.0)


# =================================================================================================
# lamb
# =================================================================================================
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("wd,nvlamb", [(0.0, False), (0.01, False), (0.0, True)])
@pytest.mark.parametrize("max_norm", [1.0, 1e9])
def test_lamb_fp32_matches_ref_and_oracle(amp_C, ref, mode, wd, nvlamb, max_norm):
    lists = _adam_state(torch.float32, seed=30, sizes=SIZES + [200003])
    r, c = clone(lists), cpu(lists)
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    gn, _ = amp_C.multi_tensor_l2norm(CHUNK, f, [lists[0]], False)
    args = (1e-2, 0.9, 0.999, 1e-6, 4, 1, wd, 1, mode)
    amp_C.multi_tensor_lamb(CHUNK, f, lists, *args, gn, max_norm, nvlamb)
    ref.multi_tensor_lamb(CHUNK, rf, r, *args, gn.clone(), max_norm, nvlamb)
    O.multi_tensor_lamb(CHUNK, of, c, *args, float(gn), max_norm, nvlamb)
    assert int(f) == int(rf) == of[0] == 0
    # update (left in g), m, v: element-wise math, 1e-6; p carries the trust ratio (norms 1e-5)
    for k, rt in ((0, 2e-6), (2, 1e-6), (1, 1e-5)):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=rt, scale="self", what="list %d vs ref" % k)
            assert_close_rel(a, cc, rtol=rt, scale="self", what="list %d vs oracle" % k)
    assert_rel_pure(lists[3], r[3], 2e-6, "v vs ref")                                        # no floor for v
    assert_rel_pure(lists[3], [x.cuda() for x in c[3]], 2e-6, "v vs oracle")


@pytest.mark.parametrize("dt", ["f16", "bf16"])
def test_lamb_16bit(amp_C, ref, dt):
    lists = _adam_state(DT[dt], seed=31)
    r = clone(lists)
    gn, _ = amp_C.multi_tensor_l2norm(CHUNK, flag_tensor(), [lists[0]], False)
    args = (1e-2, 0.9, 0.999, 1e-6, 2, 1, 0.01, 1, 1)
    amp_C.multi_tensor_lamb(CHUNK, flag_tensor(), lists, *args, gn, 1.0, True)
    ref.multi_tensor_lamb(CHUNK, flag_tensor(), r, *args, gn.clone(), 1.0, True)
    ulp = {"f16": 2e-3, "bf16": 1.6e-2}[dt]
    for k in range(4):
        for a, b in zip(lists[k], r[k]):
            assert_close_rel(a, b, rtol=ulp, scale="self")


@pytest.mark.parametrize("depth", [4, 5])
@pytest.mark.parametrize("nvlamb", [False, True])
def test_lamb_mp_matches_ref_and_oracle(amp_C, ref, depth, nvlamb):
    gdt = torch.float16 if depth == 5 else torch.float32
    (g,) = rand_lists(SIZES, [gdt], seed=32, scale=100.0)
    _, p, m, v = _adam_state(torch.float32, seed=33)
    lists = [g, p, m, v]
    if depth == 5:
        lists.append([x.to(torch.float16) for x in p])
    r, c = clone(lists), cpu(lists)
    dev = "cuda"
    lr = torch.tensor([2e-3], device=dev)
    step = torch.tensor([3], device=dev, dtype=torch.int32)
    scale = 128.0
    inv_scale = torch.tensor([1.0 / scale], device=dev)
    found_inf = torch.zeros(1, device=dev)
    max_norm = torch.tensor([1.0 * scale], device=dev)
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    gn, _ = amp_C.multi_tensor_l2norm_mp(CHUNK, f, [g], False)
    rgn, _ = ref.multi_tensor_l2norm_mp(CHUNK, rf, [r[0]], False)
    assert_close_rel(gn, rgn, rtol=1e-5)
    args = (0.9, 0.999, 1e-6)
    amp_C.multi_tensor_lamb_mp(CHUNK, f, lists, lr, *args, step, 1, 0.01, 1, 1, gn, max_norm, nvlamb, found_inf, inv_scale)
    ref.multi_tensor_lamb_mp(CHUNK, rf, r, lr, *args, step, 1, 0.01, 1, 1, gn.clone(), max_norm, nvlamb, found_inf, inv_scale)
    O.multi_tensor_lamb_mp(CHUNK, of, c, float(lr), *args, int(step), 1, 0.01, 1, 1, float(gn), float(max_norm), nvlamb, 1.0 / scale)
    assert int(f) == int(rf) == of[0] == 0
    g_rt = 1e-3 if depth == 5 else 2e-6
    for k, rt in ((0, g_rt), (2, 1e-6), (1, 1e-5)):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=rt, scale="self", what="list %d vs ref" % k)
            assert_close_rel(a, cc, rtol=rt, scale="self", what="list %d vs oracle" % k)
    assert_rel_pure(lists[3], r[3], 2e-6, "v vs ref")
    assert_rel_pure(lists[3], [x.cuda() for x in c[3]], 2e-6, "v vs oracle")
    if depth == 5:
        for a, b in zip(lists[4], r[4]):
            assert_close_rel(a, b, rtol=1e-3, scale="self", what="fp16 model copy")


def test_lamb_mp_noop_skips_everything(amp_C):
    (g,) = rand_lists(SIZES, [torch.float16], seed=34)
    _, p, m, v = _adam_state(torch.float32, seed=35)
    lists = [g, p, m, v, [x.to(torch.float16) for x in p]]
    before = clone(lists)
    one = torch.ones(1, device="cuda")
    amp_C.multi_tensor_lamb_mp(CHUNK, flag_tensor(1), lists, one * 1e-3, 0.9, 0.999, 1e-6,
                               torch.ones(1, device="cuda", dtype=torch.int32), 1, 0.01, 1, 1, one, one, True, one * 0, one)
    for l0, l1 in zip(before, lists):
        for a, b in zip(l0, l1):
            assert torch.equal(a, b)


def test_lamb_split_stages(amp_C, ref):
    sizes = [5, 4099, 70001]
    g, p, m, v = _adam_state(torch.float32, seed=36, sizes=sizes)
    upd = [torch.zeros_like(x) for x in p]
    lists = [g, p, m, v, upd]
    r = clone(lists)
    decay = torch.tensor([0.0, 0.01, 0.1], device="cuda")
    gnorm = torch.tensor([3.0])
    amp_C.multi_tensor_lamb_stage1_cuda(CHUNK, flag_tensor(), lists, decay, 4, 0.9, 0.999, 1e-6, gnorm, 1.0)
    ref.multi_tensor_lamb_stage1_cuda(CHUNK, flag_tensor(), r, decay, 4, 0.9, 0.999, 1e-6, gnorm, 1.0)
    for k in (2, 3, 4):
        for a, b in zip(lists[k], r[k]):
            assert_close_rel(a, b, rtol=2e-6, scale="self")
    pn = torch.stack([x.norm() for x in p])
    un = torch.stack([x.norm() for x in upd])
    amp_C.multi_tensor_lamb_stage2_cuda(CHUNK, flag_tensor(), [p, upd], pn, un, 1e-2, 0.01, False)
    ref.multi_tensor_lamb_stage2_cuda(CHUNK, flag_tensor(), [r[1], r[4]], pn, un, 1e-2, 0.01, False)
    for a, b in zip(p, r[1]):
        assert_close_rel(a, b, rtol=1e-6, scale="self")


# =================================================================================================
# sharded-path kernels: distributed_adam_cuda, fused_adam_cuda
# =================================================================================================
@pytest.mark.parametrize("gdt", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("depth", [4, 5])
@pytest.mark.parametrize("mode", [0, 1])
def test_dist_adam_matches_ref_and_oracle(gdt, depth, mode):
    import distributed_adam_cuda as ours
    refm = load_ref("distributed_adam_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/distributed_adam_cuda_ref.so not built")
    sizes = SIZES
    T = len(sizes)
    _, p, m, v = _adam_state(torch.float32, seed=40)
    (g,) = rand_lists(sizes, [DT[gdt]], seed=41, scale=32.0)
    lists = [p, m, v, g]
    if depth == 5:
        lists.append([torch.zeros(n, device="cuda", dtype=DT[gdt]) for n in sizes])
    r, c = clone(lists), cpu(lists)
    b1 = torch.linspace(0.85, 0.95, T, device="cuda")
    b2 = torch.linspace(0.99, 0.999, T, device="cuda")
    bc = torch.tensor([1, 0] * (T // 2) + [1] * (T % 2), device="cuda", dtype=torch.int32)
    eps = torch.full((T,), 1e-8, device="cuda")
    wd = torch.linspace(0.0, 0.1, T, device="cuda")
    args = (1e-3, 32.0, 5, mode)
    ours.multi_tensor_fused_adam(CHUNK, flag_tensor(), lists, b1, b2, bc, eps, wd, *args)
    refm.multi_tensor_fused_adam(CHUNK, flag_tensor(), r, b1, b2, bc, eps, wd, *args)
    O.multi_tensor_fused_adam(c, b1.tolist(), b2.tolist(), bc.tolist(), eps.tolist(), wd.tolist(), *args)
    for k in (0, 1):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=1e-6, scale="self", what="list %d vs ref" % k)
            assert_close_rel(a, cc, rtol=1e-6, scale="self", what="list %d vs oracle" % k)
    assert_rel_pure(lists[2], r[2], 1e-6, "v vs ref")
    assert_rel_pure(lists[2], [x.cuda() for x in c[2]], 1e-6, "v vs oracle")
    if depth == 5:
        ulp = {"f16": 1e-3, "bf16": 8e-3, "f32": 1e-6}[gdt]
        for a, b in zip(lists[4], r[4]):
            assert_close_rel(a, b, rtol=ulp, scale="self")
        for a, b in zip(lists[4], lists[0]):
            assert torch.equal(a, b.to(DT[gdt]))   # the copy is exactly the rounded master weight


@pytest.mark.parametrize("depth", [4, 5])
def test_contrib_adam_mt_matches_ref(depth):
    import fused_adam_cuda as ours
    refm = load_ref("fused_adam_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/fused_adam_cuda_ref.so not built")
    _, p, m, v = _adam_state(torch.float32, seed=42)
    (g,) = rand_lists(SIZES, [torch.float16], seed=43, scale=32.0)
    lists = [p, m, v, g]
    if depth == 5:
        lists.append([torch.zeros(n, device="cuda", dtype=torch.float16) for n in SIZES])
    r, c = clone(lists), cpu(lists)
    args = (1e-3, 0.9, 0.999, 1e-8, 32.0, 4, 1, 1, 0.01)
    ours.adam_mt(CHUNK, flag_tensor(), lists, *args)
    refm.adam_mt(CHUNK, flag_tensor(), r, *args)
    O.adam_mt(c, *args)
    for k in (0, 1):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=1e-6, scale="self", what="vs ref")
            assert_close_rel(a, cc, rtol=1e-6, scale="self", what="vs oracle")
    assert_rel_pure(lists[2], r[2], 1e-6, "v vs ref")
    assert_rel_pure(lists[2], [x.cuda() for x in c[2]], 1e-6, "v vs oracle")


def test_maybe_cast_and_strided_check(amp_C):
    import fused_adam_cuda as ours
    refm = load_ref("fused_adam_cuda_ref")
    ins, outs = rand_lists(SIZES, [torch.float16, torch.float32], seed=44)
    ours.maybe_cast_mt(CHUNK, flag_tensor(0), [ins, outs])
    for a, b in zip(ins, outs):
        assert torch.equal(a.float(), b)
    keep = [o.clone() for o in outs]
    ours.maybe_cast_mt(CHUNK, flag_tensor(1), [[x * 2 for x in ins], outs])
    for a, b in zip(keep, outs):
        assert torch.equal(a, b)                # skipped on overflow
    x = torch.ones(100000, device="cuda", dtype=torch.float16)
    for idx, stride, clear, start, expect in [(64, 64, 1, 1, 1), (65, 64, 1, 1, 0), (65, 64, 0, 1, 1), (99968, 32, 1, 0, 1), (5, 1, 0, 0, 1)]:
        y = x.clone()
        y[idx] = float("inf")
        f = flag_tensor(start)
        ours.strided_check_finite(f, y, stride, clear)
        assert int(f) == expect, (idx, stride, clear, start)
        if refm is not None:
            rf = flag_tensor(start)
            refm.strided_check_finite(rf, y, stride, clear)
            assert int(rf) == expect
