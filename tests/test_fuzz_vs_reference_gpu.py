"""GPU: seeded differential fuzz of the drop-in ops against the reference's own kernels (oracle/_ref/amp_C_ref.so, built
unmodified for sm_100a by oracle/build_ref.sh).  Where tests/test_parity_gpu.py walks a fixed list of sizes through every
dtype combination, this file draws the *shape of the call* at random per seed: how many tensors (past the reference's
110/64/48/36-per-launch caps, csrc/multi_tensor_apply.cuh:16-26), sizes sitting on and around the tile / vector / chunk
boundaries, every tensor a view at a random element offset into one arena (so vector alignment differs per tensor and
per list), a random chunk_size, random hyper-parameters and -- for the ops that raise the overflow flag -- a non-finite value at a random
place.  Same tolerances as test_parity_gpu.py (north_star): flags and scale/cast results bit-exact, fp32 Adam/SGD 1e-6,
norms and LAMB 1e-5, 16-bit storage one ulp of the storage type."""
import random

import pytest
import torch

from util import assert_close_rel, assert_rel_pure, load_ref

pytestmark = pytest.mark.gpu

DTS = [torch.float32, torch.float16, torch.bfloat16]
ULP = {torch.float32: 2e-7, torch.float16: 1e-3, torch.bfloat16: 8e-3}
EDGE_SIZES = [1, 2, 3, 5, 7, 8, 9, 31, 32, 33, 255, 256, 257, 1023, 1024, 1025, 4095, 4096, 4097, 8191, 8192, 8193,
              12289, 16384, 32767, 65535, 65536, 65537, 100001, 131072]
CHUNKS = [64, 333, 2048, 4096, 33333, 2048 * 32]
SEEDS = list(range(24))


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


def flag():
    return torch.zeros(1, dtype=torch.int32, device="cuda")


class Case(object):
    """The random shape of one call: tensor count, sizes, chunk size, and a generator for arena-backed lists."""

    def __init__(self, seed, max_tensors=130):
        self.rng = random.Random(1000 + seed)
        self.gen = torch.Generator(device="cuda").manual_seed(seed)
        r = self.rng
        self.T = r.choice([1, 2, 3, 7, 37, 49, 65, 111, max_tensors])
        big = 1 if self.T > 40 else 4               # keep the arena of a many-tensor case small
        self.sizes = [r.choice(EDGE_SIZES) if r.random() < 0.7 else r.randrange(1, 200000 // (5 - big)) for _ in range(self.T)]
        self.chunk = r.choice(CHUNKS)
        if self.chunk == 64:                        # the reference relaunches every 320 chunks: keep that case quick
            self.sizes = [min(n, 5000) for n in self.sizes]

    def views(self, dtype, scale=1.0, positive=False):
        """One list: every tensor a view into a single arena at a random 0..7-element misalignment."""
        offs, total = [], 0
        for n in self.sizes:
            total += self.rng.randrange(0, 8)
            offs.append(total)
            total += n
        arena = torch.randn(total + 8, generator=self.gen, device="cuda") * scale
        if positive:
            arena = arena.abs()
        arena = arena.to(dtype)
        return [arena[o:o + n] for o, n in zip(offs, self.sizes)]

    def poison(self, lst):
        """Put one inf / -inf / nan somewhere in `lst`; -> description."""
        t = self.rng.randrange(len(lst))
        i = self.rng.randrange(lst[t].numel())
        v = self.rng.choice([float("inf"), -float("inf"), float("nan")])
        lst[t][i] = v
        return "tensor %d index %d value %r" % (t, i, v)


def clone(lists):
    return [[t.clone() for t in l] for l in lists]


def same_bits(a, b):
    """Bit-for-bit except that any NaN matches any NaN."""
    return torch.equal(torch.nan_to_num(a.float(), nan=12345.0), torch.nan_to_num(b.float(), nan=12345.0))


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_scale(amp_C, ref, seed):
    c = Case(seed)
    r = c.rng
    in_t, out_t = r.choice(DTS), r.choice(DTS)
    xs, outs = c.views(in_t, scale=r.choice([1.0, 100.0, 30000.0])), c.views(out_t)
    where = c.poison(xs) if r.random() < 0.4 else None
    r_outs = [o.clone() for o in outs]
    s = r.choice([1.0 / 128.0, 0.5, 3.0, 1.0 / 65536.0])
    f, rf = flag(), flag()
    amp_C.multi_tensor_scale(c.chunk, f, [xs, outs], s)
    ref.multi_tensor_scale(c.chunk, rf, [xs, r_outs], s)
    expect = int(any(not bool(torch.isfinite(x).all()) for x in xs))      # 30000 * randn leaves fp16's range now and then
    assert where is None or expect == 1
    assert int(f) == int(rf) == expect, (where, c.T, c.chunk)
    for k, (a, b) in enumerate(zip(outs, r_outs)):
        assert same_bits(a, b), "tensor %d of %d, chunk %d, %s -> %s" % (k, c.T, c.chunk, in_t, out_t)


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_axpby(amp_C, ref, seed):
    c = Case(seed, max_tensors=90)
    r = c.rng
    x_t, y_t, o_t = r.choice(DTS), r.choice(DTS), r.choice(DTS)
    xs, ys, outs = c.views(x_t, scale=10.0), c.views(y_t, scale=10.0), c.views(o_t)
    arg = r.choice([-1, 0, 1])
    which = r.choice([None, None, 0, 1])
    if which is not None:
        c.poison(xs if which == 0 else ys)
    expect = 0 if which is None else int(arg == -1 or arg == which)
    assert all(bool(torch.isfinite(t).all()) for l, w in ((xs, 0), (ys, 1)) if w != which for t in l)
    rl = [xs, ys, [o.clone() for o in outs]]
    a_, b_ = r.choice([2.0, 1.0, -0.5]), r.choice([0.37, 1.0, 0.0])
    f, rf = flag(), flag()
    amp_C.multi_tensor_axpby(c.chunk, f, [xs, ys, outs], a_, b_, arg)
    ref.multi_tensor_axpby(c.chunk, rf, rl, a_, b_, arg)
    assert int(f) == int(rf) == expect, (which, arg)
    for k, (a, b) in enumerate(zip(outs, rl[2])):
        assert_close_rel(a, b, rtol=ULP[o_t], scale="self", what="tensor %d" % k)


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_l2norm(amp_C, ref, seed):
    c = Case(seed)
    r = c.rng
    dt = r.choice(DTS)
    xs = c.views(dt, scale=r.choice([1e-3, 1.0, 50.0]))
    per_tensor = r.random() < 0.6
    poisoned = r.random() < 0.25
    if poisoned:
        c.poison(xs)
    f, rf = flag(), flag()
    n, p = amp_C.multi_tensor_l2norm(c.chunk, f, [xs], per_tensor)
    rn, rp = ref.multi_tensor_l2norm(c.chunk, rf, [xs], per_tensor)
    assert int(f) == int(rf) == int(poisoned)
    assert n.shape == rn.shape and p.shape == rp.shape
    if not poisoned:
        assert_rel_pure(n, rn, 1e-5, "norm (T=%d chunk=%d)" % (c.T, c.chunk))
        if per_tensor:
            assert_rel_pure(p, rp, 1e-5, "per-tensor norms")
    else:
        assert not bool(torch.isfinite(n).all()) and not bool(torch.isfinite(rn).all())


def adam_state(c, dt):
    g, p = c.views(dt), c.views(dt)
    m, v = c.views(dt, scale=0.1), c.views(dt, scale=0.01, positive=True)
    return [g, p, m, v]


def operand_scale(lists):
    """Per tensor: the largest magnitude among g, p and m -- what one ulp of a first-moment update is measured against."""
    g, p, m = lists[0], lists[1], lists[2]
    mx = torch.stack([torch.stack([x.float().abs().max() for x in l]) for l in (g, p, m)]).max(0).values
    return mx.tolist()


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_adam(amp_C, ref, seed):
    c = Case(seed, max_tensors=120)
    r = c.rng
    dt = r.choice([torch.float32, torch.float32, torch.float16, torch.bfloat16])
    lists = adam_state(c, dt)
    rl = clone(lists)
    f = flag()
    rt = 1e-6 if dt == torch.float32 else ULP[dt]
    for step in sorted(r.sample(range(1, 2000), 2)):
        args = (r.choice([1e-4, 1e-3, 1e-2]), r.choice([0.0, 0.8, 0.9]), r.choice([0.95, 0.999]), r.choice([1e-8, 1e-6]),
                step, r.choice([0, 1]), r.choice([0, 1]), r.choice([0.0, 0.01, 0.1]))
        opnd = operand_scale(lists)
        amp_C.multi_tensor_adam(c.chunk, f, lists, *args)
        ref.multi_tensor_adam(c.chunk, f, rl, *args)
        for i, (a, b) in enumerate(zip(lists[1], rl[1])):
            assert_close_rel(a, b, rtol=rt, scale="self", what="p[%d] step %d args %r" % (i, step, args))
        for i, (a, b) in enumerate(zip(lists[2], rl[2])):   # m = b1*m + (1-b1)*g may cancel: an ulp of the operands, not of m
            assert_close_rel(a, b, rtol=rt, scale=opnd[i], what="m[%d] step %d args %r" % (i, step, args))
        if dt == torch.float32:
            assert_rel_pure(lists[3], rl[3], 1e-6, "v step %d" % step)
        else:
            for a, b in zip(lists[3], rl[3]):
                assert_close_rel(a, b, rtol=rt, scale="self", what="v step %d" % step)
    assert int(f) == 0
    for a, b in zip(lists[0], rl[0]):
        assert torch.equal(a, b)                    # gradients are never written


SGD_CASES = [(torch.float16, torch.float16, 3), (torch.float32, torch.float32, 3), (torch.float16, torch.float32, 4),
             (torch.float32, torch.float32, 4), (torch.bfloat16, torch.bfloat16, 3), (torch.bfloat16, torch.float32, 4)]


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_sgd(amp_C, ref, seed):
    c = Case(seed, max_tensors=120)
    r = c.rng
    gdt, wdt, depth = r.choice(SGD_CASES)
    lists = [c.views(gdt, scale=64.0), c.views(wdt), c.views(wdt)]
    if depth == 4:
        lists.append([torch.zeros(n, device="cuda", dtype=torch.float16) for n in c.sizes])
    rl = clone(lists)
    momentum = r.choice([0.0, 0.5, 0.9])
    nesterov = momentum > 0 and r.random() < 0.5
    dampening = 0.0 if nesterov else r.choice([0.0, 0.1])
    wd, wd_after, lr = r.choice([0.0, 1e-4, 1e-2]), r.random() < 0.5, r.choice([0.1, 0.01])
    rt = 1e-6 if wdt == torch.float32 else ULP[wdt]
    for first_run in (True, False):
        args = (wd, momentum, dampening, lr, nesterov, first_run, wd_after, 1.0 / 64.0)
        f, rf = flag(), flag()
        amp_C.multi_tensor_sgd(c.chunk, f, lists, *args)
        ref.multi_tensor_sgd(c.chunk, rf, rl, *args)
        assert int(f) == int(rf) == 0
        for k in range(1, depth):
            for i, (a, b) in enumerate(zip(lists[k], rl[k])):
                assert_close_rel(a, b, rtol=1e-3 if k == 3 else rt, scale="self", what="list %d[%d] args %r" % (k, i, args))


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_lamb(amp_C, ref, seed):
    c = Case(seed, max_tensors=100)
    r = c.rng
    lists = adam_state(c, torch.float32)
    rl = clone(lists)
    f, rf = flag(), flag()
    gn, _ = amp_C.multi_tensor_l2norm(c.chunk, f, [lists[0]], False)
    args = (r.choice([1e-3, 1e-2]), r.choice([0.8, 0.9]), r.choice([0.95, 0.999]), 1e-6, r.randrange(1, 500), r.choice([0, 1]),
            r.choice([0.0, 0.01]), r.choice([0, 1]), r.choice([0, 1]))
    max_norm, nvlamb = r.choice([0.1, 1.0, 1e9]), r.random() < 0.5
    opnd = operand_scale(lists)
    amp_C.multi_tensor_lamb(c.chunk, f, lists, *args, gn, max_norm, nvlamb)
    ref.multi_tensor_lamb(c.chunk, rf, rl, *args, gn.clone(), max_norm, nvlamb)
    assert int(f) == int(rf) == 0
    for k, rt in ((0, 2e-6), (1, 1e-5)):                 # update (left in g): element-wise; p carries the trust ratio
        for i, (a, b) in enumerate(zip(lists[k], rl[k])):
            assert_close_rel(a, b, rtol=rt, scale="self", what="list %d[%d] args %r" % (k, i, args))
    for i, (a, b) in enumerate(zip(lists[2], rl[2])):
        assert_close_rel(a, b, rtol=1e-6, scale=opnd[i], what="m[%d] args %r" % (i, args))
    assert_rel_pure(lists[3], rl[3], 2e-6, "v")


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_l2norm_scale(amp_C, ref, seed):
    c = Case(seed, max_tensors=100)
    r = c.rng
    in_t, out_t = r.choice([(torch.float32, torch.float32), (torch.float16, torch.float32), (torch.float32, torch.float16),
                            (torch.float16, torch.float16)])   # the reference dispatches fp32 / fp16 only here
    xs, outs = c.views(in_t, scale=r.choice([1.0, 20.0])), c.views(out_t)
    r_outs = [o.clone() for o in outs]
    per_tensor = r.random() < 0.6
    s = r.choice([0.125, 1.0 / 1024.0, 2.0])
    f, rf = flag(), flag()
    n, p = amp_C.multi_tensor_l2norm_scale(c.chunk, f, [xs, outs], s, per_tensor)
    rn, rp = ref.multi_tensor_l2norm_scale(c.chunk, rf, [xs, r_outs], s, per_tensor)
    assert int(f) == int(rf) == 0
    assert_rel_pure(n, rn, 1e-5, "norm")
    if per_tensor:
        assert_rel_pure(p, rp, 1e-5, "per-tensor norms")
    for k, (a, b) in enumerate(zip(outs, r_outs)):
        assert same_bits(a, b), "tensor %d" % k


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_lamb_mp(amp_C, ref, seed):
    c = Case(seed, max_tensors=100)
    r = c.rng
    depth = r.choice([4, 5])
    gdt = torch.float16 if depth == 5 else torch.float32
    scale = r.choice([1.0, 128.0, 1024.0])
    g = c.views(gdt, scale=scale if gdt == torch.float32 else min(scale, 128.0))
    _, p, m, v = adam_state(c, torch.float32)
    lists = [g, p, m, v]
    if depth == 5:
        lists.append([x.to(torch.float16) for x in p])
    rl = clone(lists)
    lr = torch.tensor([r.choice([2e-3, 1e-2])], device="cuda")
    step = torch.tensor([r.randrange(1, 300)], device="cuda", dtype=torch.int32)
    inv_scale = torch.tensor([1.0 / scale], device="cuda")
    found_inf = torch.zeros(1, device="cuda")
    max_norm = torch.tensor([r.choice([0.5, 1.0, 1e6]) * scale], device="cuda")
    f, rf = flag(), flag()
    gn, _ = amp_C.multi_tensor_l2norm_mp(c.chunk, f, [g], False)
    rgn, _ = ref.multi_tensor_l2norm_mp(c.chunk, rf, [rl[0]], False)
    assert_rel_pure(gn, rgn, 1e-5, "gradient norm")
    args = (r.choice([0.8, 0.9]), r.choice([0.95, 0.999]), 1e-6)
    tail = (step, r.choice([0, 1]), r.choice([0.0, 0.01]), r.choice([0, 1]), r.choice([0, 1]))
    nvlamb = r.random() < 0.5
    opnd = operand_scale([[x.float() / scale for x in g], p, m])
    amp_C.multi_tensor_lamb_mp(c.chunk, f, lists, lr, *args, *tail, gn, max_norm, nvlamb, found_inf, inv_scale)
    ref.multi_tensor_lamb_mp(c.chunk, rf, rl, lr, *args, *tail, gn.clone(), max_norm, nvlamb, found_inf, inv_scale)
    assert int(f) == int(rf) == 0
    what = "depth %d args %r %r nvlamb %r" % (depth, args, tail[1:], nvlamb)
    for k, rt in ((0, 1e-3 if depth == 5 else 2e-6), (1, 1e-5)):
        for i, (a, b) in enumerate(zip(lists[k], rl[k])):
            assert_close_rel(a, b, rtol=rt, scale="self", what="list %d[%d] %s" % (k, i, what))
    for i, (a, b) in enumerate(zip(lists[2], rl[2])):
        assert_close_rel(a, b, rtol=1e-6, scale=opnd[i], what="m[%d] %s" % (i, what))
    assert_rel_pure(lists[3], rl[3], 2e-6, "v")
    if depth == 5:
        for a, b in zip(lists[4], rl[4]):
            assert_close_rel(a, b, rtol=1e-3, scale="self", what="fp16 model copy")


@pytest.mark.parametrize("seed", SEEDS)
def test_fuzz_sharded_adam_kernel(seed):
    """distributed_adam_cuda.multi_tensor_fused_adam ([p, m, v, g(, p_copy)], per-tensor hyper-parameter arrays)."""
    import distributed_adam_cuda as ours
    refm = load_ref("distributed_adam_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/distributed_adam_cuda_ref.so not built")
    c = Case(seed, max_tensors=100)
    r = c.rng
    gdt = r.choice(DTS)
    depth = r.choice([4, 5])
    gscale = r.choice([1.0, 32.0])
    _, p, m, v = adam_state(c, torch.float32)
    g = c.views(gdt, scale=gscale)
    lists = [p, m, v, g]
    if depth == 5:
        lists.append([torch.zeros(n, device="cuda", dtype=gdt) for n in c.sizes])
    rl = clone(lists)
    T = c.T
    rnd = lambda lo, hi: torch.tensor([r.uniform(lo, hi) for _ in range(T)], device="cuda")
    b1, b2, eps, wd = rnd(0.8, 0.95), rnd(0.95, 0.999), rnd(1e-8, 1e-6), rnd(0.0, 0.1)
    bc = torch.tensor([r.choice([0, 1]) for _ in range(T)], device="cuda", dtype=torch.int32)
    args = (r.choice([1e-4, 1e-3]), gscale, r.randrange(1, 1000), r.choice([0, 1]))
    opnd = operand_scale([[x.float() / gscale for x in g], p, m])
    ours.multi_tensor_fused_adam(c.chunk, flag(), lists, b1, b2, bc, eps, wd, *args)
    refm.multi_tensor_fused_adam(c.chunk, flag(), rl, b1, b2, bc, eps, wd, *args)
    for i, (a, b) in enumerate(zip(lists[0], rl[0])):
        assert_close_rel(a, b, rtol=1e-6, scale="self", what="p[%d] args %r" % (i, args))
    for i, (a, b) in enumerate(zip(lists[1], rl[1])):
        assert_close_rel(a, b, rtol=1e-6, scale=opnd[i], what="m[%d] args %r" % (i, args))
    assert_rel_pure(lists[2], rl[2], 1e-6, "v")
    if depth == 5:
        for a, b in zip(lists[4], lists[0]):
            assert torch.equal(a, b.to(gdt))             # the copy is exactly the rounded master weight
        for a, b in zip(lists[4], rl[4]):
            assert_close_rel(a, b, rtol=ULP[gdt] if gdt != torch.float32 else 1e-6, scale="self", what="p_copy")


@pytest.mark.parametrize("seed", SEEDS[:12])
def test_fuzz_contrib_adam_mt(seed):
    """fused_adam_cuda.adam_mt ([p, m, v, g(, p_copy)], the legacy contrib FusedAdam kernel)."""
    import fused_adam_cuda as ours
    refm = load_ref("fused_adam_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/fused_adam_cuda_ref.so not built")
    c = Case(seed, max_tensors=100)
    r = c.rng
    depth = r.choice([4, 5])
    gscale = r.choice([1.0, 32.0])
    _, p, m, v = adam_state(c, torch.float32)
    g = c.views(torch.float16, scale=gscale)
    lists = [p, m, v, g]
    if depth == 5:
        lists.append([torch.zeros(n, device="cuda", dtype=torch.float16) for n in c.sizes])
    rl = clone(lists)
    args = (r.choice([1e-4, 1e-3]), r.choice([0.8, 0.9]), r.choice([0.95, 0.999]), r.choice([1e-8, 1e-6]), gscale,
            r.randrange(1, 1000), r.choice([0, 1]), r.choice([0, 1]), r.choice([0.0, 0.01]))
    opnd = operand_scale([[x.float() / gscale for x in g], p, m])
    ours.adam_mt(c.chunk, flag(), lists, *args)
    refm.adam_mt(c.chunk, flag(), rl, *args)
    for i, (a, b) in enumerate(zip(lists[0], rl[0])):
        assert_close_rel(a, b, rtol=1e-6, scale="self", what="p[%d] args %r" % (i, args))
    for i, (a, b) in enumerate(zip(lists[1], rl[1])):
        assert_close_rel(a, b, rtol=1e-6, scale=opnd[i], what="m[%d] args %r" % (i, args))
    assert_rel_pure(lists[2], rl[2], 1e-6, "v")
    if depth == 5:
        for a, b in zip(lists[4], rl[4]):
            assert_close_rel(a, b, rtol=1e-3, scale="self", what="p_copy")
