"""GPU, at the FULL sizes BASELINE.json names (ResNet-50: 161 tensors / 25.6 M params; BERT-large: 398 tensors /
336 M params): size-independent properties that need no oracle run -- exact round trips, linearity, scale
covariance of norms, fixed points of the optimizers, idempotent flags -- so that parity is not only shown at
the small sizes the CPU oracle can finish."""
import math

import pytest
import torch

import param_shapes

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32


@pytest.fixture(scope="module")
def amp_C():
    import amp_C as m
    return m


def flag_tensor(v=0):
    return torch.full((1,), v, dtype=torch.int32, device="cuda")


def make(shapes, dtype=torch.float32, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return [(torch.randn(s, generator=g, device="cuda") * scale).to(dtype) for s in shapes]


@pytest.fixture(scope="module", params=["resnet50", "bert_large"])
def shapes(request):
    return param_shapes.MODELS[request.param]()


def test_scale_round_trip_is_exact_for_powers_of_two(amp_C, shapes):
    x = make(shapes, scale=65536.0)
    keep = [t.clone() for t in x]
    f = flag_tensor()
    amp_C.multi_tensor_scale(CHUNK, f, [x, x], 1.0 / 65536.0)      # the amp unscale, in place
    amp_C.multi_tensor_scale(CHUNK, f, [x, x], 65536.0)
    assert int(f) == 0
    assert all(torch.equal(a, b) for a, b in zip(x, keep))
    # fp32 -> fp16 -> fp32 of fp16-representable values is the identity as well (master <-> model copies)
    h = [(t / 65536.0).half() for t in keep]      # O(1) magnitudes: inside the fp16 range
    back = [torch.empty_like(t) for t in keep]
    h2 = [torch.empty_like(t) for t in h]
    amp_C.multi_tensor_scale(CHUNK, f, [h, back], 1.0)
    amp_C.multi_tensor_scale(CHUNK, f, [back, h2], 1.0)
    assert all(torch.equal(a, b) for a, b in zip(h, h2)) and int(f) == 0


def test_overflow_flag_is_found_anywhere_and_only_then(amp_C, shapes):
    x = make(shapes)
    out = [torch.empty_like(t) for t in x]
    f = flag_tensor()
    amp_C.multi_tensor_scale(CHUNK, f, [x, out], 0.5)
    assert int(f) == 0
    for ti, pos in ((0, 0), (len(x) // 2, -1), (len(x) - 1, x[-1].numel() // 2)):
        saved = x[ti].view(-1)[pos].clone()
        x[ti].view(-1)[pos] = float("inf")
        f = flag_tensor()
        amp_C.multi_tensor_scale(CHUNK, f, [x, out], 0.5)
        assert int(f) == 1, (ti, pos)
        f2 = flag_tensor()
        amp_C.multi_tensor_l2norm(CHUNK, f2, [x], False)
        assert int(f2) == 1
        x[ti].view(-1)[pos] = saved


def test_axpby_is_linear_and_matches_two_scales(amp_C, shapes):
    x, y = make(shapes, seed=1), make(shapes, seed=2)
    out = [torch.empty_like(t) for t in x]
    f = flag_tensor()
    amp_C.multi_tensor_axpby(CHUNK, f, [x, y, out], 1.0, 0.0, -1)
    assert all(torch.equal(a, b) for a, b in zip(out, x))            # a=1, b=0: copy
    amp_C.multi_tensor_axpby(CHUNK, f, [x, y, out], 0.5, 0.25, -1)   # exact: power-of-two coefficients, one FMA
    for o, a, b in zip(out[:8] + out[-8:], x[:8] + x[-8:], y[:8] + y[-8:]):
        assert torch.equal(o, torch.addcmul(0.25 * b, a, torch.tensor(0.5, device="cuda")))
    assert int(f) == 0


def test_l2norm_scale_covariance_and_decomposition(amp_C, shapes):
    x = make(shapes, seed=3)
    f = flag_tensor()
    n1, per1 = amp_C.multi_tensor_l2norm(CHUNK, f, [x], True)
    x4 = [t * 4.0 for t in x]
    n4, per4 = amp_C.multi_tensor_l2norm(CHUNK, f, [x4], True)
    assert torch.equal(n4, n1 * 4.0) and torch.equal(per4, per1 * 4.0)   # exact: every partial scales by 16
    # global^2 == sum of per-tensor^2 (1e-5), and equals torch's norm of the concatenation
    total = float(torch.sqrt((per1.double() ** 2).sum()))
    assert abs(float(n1) - total) <= 1e-5 * total
    want = math.sqrt(sum(float(t.double().pow(2).sum()) for t in x))
    assert abs(float(n1) - want) <= 1e-5 * want
    # bitwise reproducible
    n1b, per1b = amp_C.multi_tensor_l2norm(CHUNK, f, [x], True)
    assert torch.equal(n1, n1b) and torch.equal(per1, per1b) and int(f) == 0


def test_adam_fixed_points(amp_C, shapes):
    p = make(shapes, scale=0.02, seed=4)
    keep = [t.clone() for t in p]
    g = [torch.zeros_like(t) for t in p]
    m = [torch.zeros_like(t) for t in p]
    v = [torch.zeros_like(t) for t in p]
    f = flag_tensor()
    # zero gradient, zero state, no decay: nothing moves, exactly
    amp_C.multi_tensor_adam(CHUNK, f, [g, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 1, 1, 1, 0.0)
    assert all(torch.equal(a, b) for a, b in zip(p, keep))
    assert all(float(t.abs().max()) == 0.0 for t in m + v)
    # first step of Adam with bias correction moves every element with g != 0 by ~lr*sign(g)
    g = make(shapes, seed=5)
    amp_C.multi_tensor_adam(CHUNK, f, [g, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 1, 1, 1, 0.0)
    for a, b, gg in zip(p[:6] + p[-6:], keep[:6] + keep[-6:], g[:6] + g[-6:]):
        big = gg.abs() > 1e-3
        assert torch.allclose((b - a)[big], 1e-3 * torch.sign(gg)[big], rtol=1e-4, atol=1e-9)
    # decoupled weight decay alone: p <- p*(1 - lr*wd)
    p2 = [t.clone() for t in keep]
    z = [torch.zeros_like(t) for t in p2]
    amp_C.multi_tensor_adam(CHUNK, f, [z, p2, [torch.zeros_like(t) for t in p2], [torch.zeros_like(t) for t in p2]],
                            1e-2, 0.9, 0.999, 1e-8, 1, 1, 1, 0.5)
    for a, b in zip(p2[:6] + p2[-6:], keep[:6] + keep[-6:]):
        assert torch.allclose(a, b - 1e-2 * 0.5 * b, rtol=1e-6, atol=1e-12)


def test_sgd_noop_and_plain_step(amp_C, shapes):
    p = make(shapes, scale=0.02, seed=6)
    keep = [t.clone() for t in p]
    g = make(shapes, seed=7)
    mom = [torch.zeros_like(t) for t in p]
    amp_C.multi_tensor_sgd(CHUNK, flag_tensor(1), [g, p, mom], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0)
    assert all(torch.equal(a, b) for a, b in zip(p, keep))          # flag set: nothing happens
    amp_C.multi_tensor_sgd(CHUNK, flag_tensor(0), [g, p, mom], 0.0, 0.0, 0.0, 0.25, False, True, False, 1.0)
    for a, b, gg in zip(p[:6] + p[-6:], keep[:6] + keep[-6:], g[:6] + g[-6:]):
        assert torch.equal(a, b + (-0.25 * gg))                      # exact: one multiply by a power of two, one add


def test_lamb_trust_ratio_invariants(amp_C, shapes):
    # with nvlamb the step of every tensor has norm exactly lr*||p|| (that is what the trust ratio is for)
    p = make(shapes, scale=0.02, seed=8)
    keep = [t.clone() for t in p]
    g = make(shapes, seed=9)
    m = [torch.zeros_like(t) for t in p]
    v = [torch.zeros_like(t) for t in p]
    f = flag_tensor()
    gn, _ = amp_C.multi_tensor_l2norm(CHUNK, f, [g], False)
    lr = 1e-2
    amp_C.multi_tensor_lamb(CHUNK, f, [g, p, m, v], lr, 0.9, 0.999, 1e-6, 1, 1, 0.0, 1, 1, gn, 1.0, True)
    assert int(f) == 0
    checked = 0
    for a, b in zip(p, keep):
        pn = float(b.double().norm())
        if b.numel() < 8 or pn == 0.0:
            continue
        step_norm = float((a.double() - b.double()).norm())
        assert abs(step_norm - lr * pn) <= 2e-4 * lr * pn + 1e-7 * pn    # p is fp32: the step is observed through rounding of p
        checked += 1
    assert checked > 100


def test_one_tensor_longer_than_2_to_the_31_elements(amp_C):
    """A single tensor of 2^31 + 4099 elements (the reference's TensorListMetadata keeps sizes in `int`,
    csrc/multi_tensor_apply.cuh:16-26, and cannot take it; here numel and the tile offsets are 64-bit): the unscale touches
    every element once, finds an inf behind the 2^31 boundary, and the norm of all-ones is sqrt(n)."""
    n = (1 << 31) + 4099
    x = torch.full((n,), 2.0, device="cuda", dtype=torch.float16)
    probes = [0, 1, 4095, 4096, (1 << 31) - 1, 1 << 31, (1 << 31) + 1, n - 2, n - 1]
    for k, i in enumerate(probes):
        x[i] = float(k + 3)
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    amp_C.multi_tensor_scale(2048 * 32, f, [[x], [x]], 0.5)
    assert int(f) == 0
    for k, i in enumerate(probes):
        assert float(x[i]) == 0.5 * (k + 3)
        x[i] = 1.0
    assert float(x.min()) == 1.0 and float(x.max()) == 1.0          # every other element was halved exactly once
    norm, per = amp_C.multi_tensor_l2norm(2048 * 32, f, [[x]], True)
    assert abs(float(norm) - n ** 0.5) <= 1e-5 * n ** 0.5 and abs(float(per[0]) - n ** 0.5) <= 1e-5 * n ** 0.5
    assert int(f) == 0
    x[(1 << 31) + 7] = float("inf")
    y = torch.empty(n, device="cuda", dtype=torch.float16)
    amp_C.multi_tensor_scale(2048 * 32, f, [[x], [y]], 1.0)
    assert int(f) == 1 and float(y[n - 1]) == 1.0 and bool(torch.isinf(y[(1 << 31) + 7]))
