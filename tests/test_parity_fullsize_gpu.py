"""GPU parity at the BASELINE.json sizes against the reference's own kernels (oracle/_ref/amp_C_ref.so = the
reference's unmodified csrc/multi_tensor_*.cu compiled for sm_100a):

  configs[1]  ResNet-50 list (161 tensors, 25,557,032 fp32 params): multi_tensor_scale, multi_tensor_adam modes 0/1, 3 steps
  configs[2]  BERT-large list (398 tensors, 336,226,108 params): multi_tensor_l2norm_mp + multi_tensor_lamb_mp on the
              5-list [g fp16, p fp32, m, v, p fp16], and multi_tensor_l2norm + multi_tensor_lamb all-fp32
  configs[4]  one 268,435,456-element tensor: scale, l2norm, axpby

This is where multi-wave grids, 31 M-element tensors (477 chunks of one tensor) and the reference's relaunch logic show up;
the small-size grids of test_parity_gpu.py cannot.

Tolerances, element-wise, in fp64:
  * scale / cast outputs and flags: bit-exact.
  * v (second moment: a sum of non-negative terms), norms, trust ratios: PURE relative, |a-b| <= rtol*|b| -- no floor.
  * m and the stored update (m = b1*m + (1-b1)*g can land near zero): rtol*|b| + rtol*max|b|.
  * p: rtol*|b| + rtol*(largest applied step) -- the error of p_new = p - lr*u is an ulp of p plus the error of the step.
    (Recorded besides, not asserted: the error against |p| + |this element's step|; where m cancels the step itself is a
    near-zero difference and that ratio says nothing.)
Each test also records the max and the 99.9th percentile of the element-wise relative error of p, m, v against the
reference; the table is written to gpurun_out/parity_errors_<library>.json (committed as profiles/parity_errors_r02.md).
"""
import json
import os

import pytest
import torch

from util import assert_close_rel, assert_rel_pure, load_ref, rel_err_stats

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLE = {}


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


@pytest.fixture(scope="module", autouse=True)
def _dump_table():
    yield
    if not TABLE:
        return
    lib = os.path.basename(os.environ.get("B200MT_LIB") or "libb200mt.so").replace(".so", "")
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_errors_%s.json" % lib), "w") as f:
        json.dump({"library": lib, "reference": "oracle/_ref/amp_C_ref.so (--use_fast_math)", "rows": TABLE}, f, indent=1)


def record(case, name, ours, want, floor=0.0):
    mx, p999, n = rel_err_stats(ours, want, floor=floor)
    TABLE.setdefault(case, {})[name] = {"max_rel": mx, "p99.9_rel": p999, "elements": n, "floor": floor}
    print("%-44s %-10s max %.3e  p99.9 %.3e  (n=%d)" % (case, name, mx, p999, n))


def record_p(case, ours, want, old):
    """p_new = p - step: the honest yardstick for its error is |p| + |applied step|, element by element."""
    worst, sample = 0.0, []
    for a, b, o in zip(ours, want, old):
        den = (b.double().abs() + (b.double() - o.double()).abs()).reshape(-1)
        r = ((a.double() - b.double()).abs().reshape(-1) / den.clamp_min(1e-300))[den > 0]
        if r.numel():
            worst = max(worst, float(r.max()))
            sample.append(r[:: max(1, r.numel() >> 16)])
    allr = torch.cat(sample)
    p999 = float(torch.kthvalue(allr, max(1, int(0.999 * allr.numel())))[0])
    TABLE.setdefault(case, {})["p_over_p_plus_step"] = {"max_rel": worst, "p99.9_rel": p999, "elements": sum(t.numel() for t in want),
                                                        "floor": "|p| + |p_new - p_old|"}
    print("%-44s %-10s max %.3e  p99.9 %.3e" % (case, "p/(p+step)", worst, p999))
    return worst


def flag():
    return torch.zeros(1, dtype=torch.int32, device="cuda")


def clone(lists):
    return [[t.clone() for t in l] for l in lists]


def make_list(shapes, seed, scale=1.0, dtype=torch.float32, positive=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    out = []
    for s in shapes:
        x = torch.randn(s, generator=g, device="cuda") * scale
        out.append((x.abs() if positive else x).to(dtype))
    return out


def upd_scale(new, old):
    return max(float((a - b).abs().max()) for a, b in zip(new, old))


# =================================================================================================
# configs[1]: ResNet-50
# =================================================================================================
def test_resnet50_scale_bit_exact(amp_C, ref):
    import param_shapes
    shapes = param_shapes.resnet50()
    g = make_list(shapes, 1, 65536.0)
    rg = [t.clone() for t in g]
    f, rf = flag(), flag()
    amp_C.multi_tensor_scale(CHUNK, f, [g, g], 1.0 / 65536.0)           # the amp unscale, in place
    ref.multi_tensor_scale(CHUNK, rf, [rg, rg], 1.0 / 65536.0)
    assert int(f) == int(rf) == 0
    assert all(torch.equal(a, b) for a, b in zip(g, rg))
    # fp16 model grads -> fp32 master grads, with one inf deep inside the largest tensor
    g16 = [t.half() for t in g]
    big = max(range(len(g16)), key=lambda i: g16[i].numel())
    g16[big].view(-1)[1234567] = float("inf")
    o, ro = [torch.empty_like(t) for t in g], [torch.empty_like(t) for t in g]
    amp_C.multi_tensor_scale(CHUNK, f, [g16, o], 0.5)
    ref.multi_tensor_scale(CHUNK, rf, [g16, ro], 0.5)
    assert int(f) == int(rf) == 1
    assert all(torch.equal(a, b) for a, b in zip(o, ro))


@pytest.mark.parametrize("mode", [0, 1])
def test_resnet50_adam_three_steps_vs_reference(amp_C, ref, mode):
    import param_shapes
    shapes = param_shapes.resnet50()
    p = make_list(shapes, 2, 0.02)
    m = [torch.zeros_like(t) for t in p]
    v = [torch.zeros_like(t) for t in p]
    ours = [None, p, m, v]
    theirs = [None] + clone([p, m, v])
    f = flag()
    for step in (1, 2, 3):
        g = make_list(shapes, 10 + step)
        ours[0], theirs[0] = g, [t.clone() for t in g]
        p_old = [t.clone() for t in theirs[1]]
        args = (1e-3, 0.9, 0.999, 1e-8, step, mode, 1, 0.01)
        amp_C.multi_tensor_adam(CHUNK, f, ours, *args)
        ref.multi_tensor_adam(CHUNK, f, theirs, *args)
        case = "resnet50 adam mode %d step %d" % (mode, step)
        record(case, "p", ours[1], theirs[1]); record(case, "m", ours[2], theirs[2]); record(case, "v", ours[3], theirs[3])
        assert_rel_pure(ours[3], theirs[3], 1e-6, "v (pure relative) " + case)
        us = upd_scale(theirs[1], p_old)
        for a, b in zip(ours[1], theirs[1]):
            assert_close_rel(a, b, rtol=1e-6, scale=us, what="p " + case)
        record_p(case, ours[1], theirs[1], p_old)      # reported only: where m cancels, the applied step is itself near zero
        for a, b in zip(ours[2], theirs[2]):
            assert_close_rel(a, b, rtol=1e-6, scale="self", what="m " + case)
        assert all(torch.equal(a, b) for a, b in zip(ours[0], theirs[0]))     # g is an input only
    assert int(f) == 0


# =================================================================================================
# configs[2]: BERT-large
# =================================================================================================
@pytest.mark.parametrize("nvlamb", [False, True])
def test_bert_large_lamb_mp_vs_reference(amp_C, ref, nvlamb):
    import param_shapes
    shapes = param_shapes.bert_large()
    scale = 1024.0
    p = make_list(shapes, 3, 0.02)
    ours = [None, p, [torch.zeros_like(t) for t in p], [torch.zeros_like(t) for t in p], [t.half() for t in p]]
    theirs = [None] + clone(ours[1:])
    lr = torch.tensor([2e-3], device="cuda")
    inv_scale = torch.tensor([1.0 / scale], device="cuda")
    found_inf = torch.zeros(1, device="cuda")
    max_norm = torch.tensor([1.0 * scale], device="cuda")
    f, rf = flag(), flag()
    for it in (1, 2):
        step = torch.tensor([it], device="cuda", dtype=torch.int32)
        g = make_list(shapes, 20 + it, scale * 0.05, torch.float16)
        ours[0], theirs[0] = g, [t.clone() for t in g]
        p_old = [t.clone() for t in theirs[1]]
        # per-tensor ||p|| through both l2norm implementations (what LAMB's trust ratio is made of)
        _, pn = amp_C.multi_tensor_l2norm(CHUNK, f, [ours[1]], True)
        _, rpn = ref.multi_tensor_l2norm(CHUNK, rf, [theirs[1]], True)
        assert_rel_pure(pn, rpn, 1e-5, "per-tensor ||p|| it %d" % it)
        gn, gpt = amp_C.multi_tensor_l2norm_mp(CHUNK, f, [ours[0]], True)
        rgn, rgpt = ref.multi_tensor_l2norm_mp(CHUNK, rf, [theirs[0]], True)
        assert_rel_pure(gn, rgn, 1e-5, "global grad norm it %d" % it)
        assert_rel_pure(gpt, rgpt, 1e-5, "per-tensor ||g|| it %d" % it)
        args = (0.9, 0.999, 1e-6)
        amp_C.multi_tensor_lamb_mp(CHUNK, f, ours, lr, *args, step, 1, 0.01, 1, 1, gn, max_norm, nvlamb, found_inf, inv_scale)
        ref.multi_tensor_lamb_mp(CHUNK, rf, theirs, lr, *args, step, 1, 0.01, 1, 1, rgn, max_norm, nvlamb, found_inf, inv_scale)
        assert int(f) == int(rf) == 0
        case = "bert-large lamb_mp nvlamb=%d step %d" % (int(nvlamb), it)
        record(case, "p", ours[1], theirs[1]); record(case, "m", ours[2], theirs[2]); record(case, "v", ours[3], theirs[3])
        record(case, "global_norm", gn, rgn); record(case, "||p||_t", pn, rpn)
        assert_rel_pure(ours[3], theirs[3], 2e-6, "v (pure relative) " + case)
        # per-tensor ||u|| over the stored update: the other half of the trust ratio
        _, un = amp_C.multi_tensor_l2norm(CHUNK, f, [ours[0]], True)
        _, run = ref.multi_tensor_l2norm(CHUNK, rf, [theirs[0]], True)
        assert_rel_pure(un, run, 1e-5, "per-tensor ||u|| " + case)
        record(case, "||u||_t", un, run)
        us = upd_scale(theirs[1], p_old)
        record_p(case, ours[1], theirs[1], p_old)
        # The update is STORED in g's dtype (fp16) between the two stages (csrc/multi_tensor_lamb_mp.cu:163-166): where the
        # two fp32 updates straddle an fp16 rounding boundary the stored values differ by one fp16 ulp (2^-10 of the
        # update), and p inherits that.  So: every element within 1e-5*|p| + 2^-10 of the largest step, and all but a
        # 1e-4 fraction of the elements within the trust-ratio tolerance proper, 1e-5*(|p| + largest step).
        loose = tight = total = 0
        for a, b in zip(ours[1], theirs[1]):
            d = (a.double() - b.double()).abs()
            loose += int((d > 1e-5 * b.double().abs() + 2.0 ** -10 * us).sum())
            tight += int((d > 1e-5 * b.double().abs() + 1e-5 * us).sum())
            total += b.numel()
        TABLE[case]["p_elements_one_fp16_ulp_of_update_apart"] = {"count": tight, "of": total}
        assert loose == 0, "p " + case
        assert tight <= 1e-4 * total, "p: %d of %d elements beyond 1e-5 (fp16 update flips) %s" % (tight, total, case)
        for a, b in zip(ours[2], theirs[2]):
            assert_close_rel(a, b, rtol=1e-6, scale="self", what="m " + case)
        for a, b in zip(ours[0], theirs[0]):
            assert_close_rel(a, b, rtol=1e-3, scale="self", what="stored update (fp16) " + case)
        for a, b in zip(ours[4], theirs[4]):       # round_fp16(p): p's own bound plus one fp16 ulp of p
            assert_close_rel(a, b, rtol=2e-3, scale=us, what="fp16 model copy " + case)


def test_bert_large_lamb_fp32_vs_reference(amp_C, ref):
    import param_shapes
    shapes = param_shapes.bert_large()
    p = make_list(shapes, 4, 0.02)
    ours = [None, p, make_list(shapes, 5, 0.01), make_list(shapes, 6, 0.001, positive=True)]
    theirs = [None] + clone(ours[1:])
    f, rf = flag(), flag()
    for it in (1, 2):
        g = make_list(shapes, 30 + it, 0.05)
        ours[0], theirs[0] = g, [t.clone() for t in g]
        p_old = [t.clone() for t in theirs[1]]
        gn, _ = amp_C.multi_tensor_l2norm(CHUNK, f, [ours[0]], False)
        rgn, _ = ref.multi_tensor_l2norm(CHUNK, rf, [theirs[0]], False)
        assert_rel_pure(gn, rgn, 1e-5, "global grad norm")
        args = (1e-2, 0.9, 0.999, 1e-6, it + 3, 1, 0.01, 1, 1)
        amp_C.multi_tensor_lamb(CHUNK, f, ours, *args, gn, 1.0, False)
        ref.multi_tensor_lamb(CHUNK, rf, theirs, *args, rgn, 1.0, False)
        assert int(f) == int(rf) == 0
        case = "bert-large lamb fp32 step %d" % it
        record(case, "p", ours[1], theirs[1]); record(case, "m", ours[2], theirs[2]); record(case, "v", ours[3], theirs[3])
        record(case, "u", ours[0], theirs[0])
        assert_rel_pure(ours[3], theirs[3], 2e-6, "v (pure relative) " + case)
        us = upd_scale(theirs[1], p_old)
        for a, b in zip(ours[1], theirs[1]):
            assert_close_rel(a, b, rtol=1e-5, scale=us, what="p " + case)
        record_p(case, ours[1], theirs[1], p_old)
        for k, name in ((2, "m"), (0, "u")):
            for a, b in zip(ours[k], theirs[k]):
                assert_close_rel(a, b, rtol=2e-6, scale="self", what=name + " " + case)


# =================================================================================================
# configs[4]: one tensor of 2^28 elements
# =================================================================================================
def test_256M_tensor_scale_l2norm_axpby_vs_reference(amp_C, ref):
    n = 1 << 28
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(n, generator=g, device="cuda")
    f, rf = flag(), flag()
    # l2norm, global and per-tensor (4096 chunks of one tensor)
    a, ap = amp_C.multi_tensor_l2norm(CHUNK, f, [[x]], True)
    b, bp = ref.multi_tensor_l2norm(CHUNK, rf, [[x]], True)
    assert_rel_pure(a, b, 1e-5, "l2norm 2^28")
    assert_rel_pure(ap, bp, 1e-5, "per-tensor l2norm 2^28")
    record("256M tensor", "l2norm", a, b)
    want = float(torch.linalg.vector_norm(x.double()))
    assert abs(float(a) - want) <= 1e-5 * want
    # scale fp32 -> fp16 and fp16 -> fp32: bit-exact
    h, rh = torch.empty(n, dtype=torch.float16, device="cuda"), torch.empty(n, dtype=torch.float16, device="cuda")
    amp_C.multi_tensor_scale(CHUNK, f, [[x], [h]], 0.37)
    ref.multi_tensor_scale(CHUNK, rf, [[x], [rh]], 0.37)
    assert torch.equal(h, rh)
    o, ro = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    amp_C.multi_tensor_scale(CHUNK, f, [[h], [o]], 3.0)
    ref.multi_tensor_scale(CHUNK, rf, [[rh], [ro]], 3.0)
    assert torch.equal(o, ro)
    del rh, ro
    # axpby: out = a*x + b*y (one FMA contraction of freedom: 1e-6 of the operands)
    y = torch.randn(n, generator=g, device="cuda")
    out, rout = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    amp_C.multi_tensor_axpby(CHUNK, f, [[x], [y], [out]], 2.0, -0.5, -1)
    ref.multi_tensor_axpby(CHUNK, rf, [[x], [y], [rout]], 2.0, -0.5, -1)
    assert int(f) == int(rf) == 0
    assert_close_rel(out, rout, rtol=1e-6, scale="self", what="axpby 2^28")
    record("256M tensor", "axpby", out, rout, floor=1e-3)
