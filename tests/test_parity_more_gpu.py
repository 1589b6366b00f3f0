"""GPU parity, continued: Adagrad and NovoGrad (amp_C.multi_tensor_adagrad / multi_tensor_novograd) against
the reference's own kernels (oracle/_ref) and the CPU oracle; FusedAdagrad / FusedNovoGrad against torch.optim
and a plain-PyTorch NovoGrad; the stand-alone NVLS kernels are covered in test_dist_adam_gpu.py."""
import numpy as np
import pytest
import torch

from oracle import mt_oracle as O
from util import assert_close_rel, load_ref

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32
SIZES = [1, 3, 555, 2048 * 32 + 1, 33333, 8192, 16, 2048 * 32 * 2 - 1]
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


def rand_lists(sizes, dtypes, seed, scale=1.0, positive=()):
    g = torch.Generator(device="cuda").manual_seed(seed)
    out = []
    for d, dt in enumerate(dtypes):
        lst = []
        for n in sizes:
            x = torch.randn(n, generator=g, device="cuda") * scale
            lst.append((x.abs() if d in positive else x).to(dt))
        out.append(lst)
    return out


def clone(lists):
    return [[t.detach().clone() for t in l] for l in lists]


def cpu(lists):
    return [[t.detach().cpu().clone() for t in l] for l in lists]


@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
@pytest.mark.parametrize("mode,wd", [(0, 0.0), (0, 0.01), (1, 0.01)])
def test_adagrad_matches_ref_and_oracle(amp_C, ref, dt, mode, wd):
    lists = rand_lists(SIZES, [DT[dt]] * 3, seed=50, positive=(2,))
    r, c = clone(lists), cpu(lists)
    for _ in range(2):
        amp_C.multi_tensor_adagrad(CHUNK, flag_tensor(), lists, 1e-2, 1e-10, mode, wd)
        ref.multi_tensor_adagrad(CHUNK, flag_tensor(), r, 1e-2, 1e-10, mode, wd)
        O.multi_tensor_adagrad(c, 1e-2, 1e-10, mode, wd)
    rt = {"f32": 1e-6, "f16": 1e-3, "bf16": 8e-3}[dt]
    for k in (1, 2):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=rt, scale="self", what="list %d vs ref" % k)
            assert_close_rel(a, cc, rtol=rt, scale="self", what="list %d vs oracle" % k)


@pytest.mark.parametrize("dt", ["f32", "f16"])
@pytest.mark.parametrize("norm_type", [0, 2])
@pytest.mark.parametrize("moment_mode", [0, 1])
def test_novograd_matches_ref_and_oracle(amp_C, ref, dt, norm_type, moment_mode):
    lists = rand_lists(SIZES, [DT[dt]] * 3, seed=51)
    r, c = clone(lists), cpu(lists)
    T = len(SIZES)
    norms = torch.linspace(0.5, 2.0, T, device="cuda")
    rnorms, onorms = norms.clone(), norms.cpu().numpy().copy()
    f, rf, of = flag_tensor(), flag_tensor(), [0]
    for step in (1, 2, 3):
        args = (1e-2, 0.95, 0.98, 1e-8, step, 1, 0.01, 1, moment_mode, norm_type)
        amp_C.multi_tensor_novograd(CHUNK, f, lists, norms, *args)
        ref.multi_tensor_novograd(CHUNK, rf, r, rnorms, *args)
        O.multi_tensor_novograd(CHUNK, of, c, onorms, *args)
        # blended per-tensor norms: 1e-5 (norm tolerance of the north star)
        assert_close_rel(norms, rnorms, rtol=1e-5, what="grad_norms vs ref step %d" % step)
        assert_close_rel(norms, torch.from_numpy(onorms), rtol=1e-5, what="grad_norms vs oracle step %d" % step)
    assert int(f) == int(rf) == of[0] == 0
    rt = {"f32": 1e-5, "f16": 2e-3}[dt]     # p and m carry the 1e-5 norm
    for k in (1, 2):
        for a, b, cc in zip(lists[k], r[k], c[k]):
            assert_close_rel(a, b, rtol=rt, scale="self", what="list %d vs ref" % k)
            assert_close_rel(a, cc, rtol=rt, scale="self", what="list %d vs oracle" % k)


def test_novograd_norm_flag_and_nan_handling(amp_C, ref):
    # L-inf: fmaxf drops NaN, inf raises the flag; L2: NaN raises the flag
    for norm_type, val, expect in [(0, float("nan"), 0), (0, float("inf"), 1), (2, float("nan"), 1), (2, float("inf"), 1)]:
        lists = rand_lists([4099, 70000], [torch.float32] * 3, seed=52)
        lists[0][1][69999] = val
        r = clone(lists)
        norms, rnorms = torch.ones(2, device="cuda"), torch.ones(2, device="cuda")
        f, rf = flag_tensor(), flag_tensor()
        args = (1e-3, 0.9, 0.99, 1e-8, 1, 1, 0.0, 1, 1, norm_type)
        amp_C.multi_tensor_novograd(CHUNK, f, lists, norms, *args)
        ref.multi_tensor_novograd(CHUNK, rf, r, rnorms, *args)
        assert int(f) == int(rf) == expect, (norm_type, val)
        assert torch.equal(torch.isfinite(norms), torch.isfinite(rnorms))


def test_fused_adagrad_vs_torch():
    from apex.optimizers import FusedAdagrad
    torch.manual_seed(9876)
    sizes = [[278011], [4096, 1024], [4096], [1]]
    ref_p = [torch.nn.Parameter(torch.rand(s, device="cuda")) for s in sizes]
    tst_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    a = torch.optim.Adagrad(ref_p, lr=5e-4, eps=1e-8, weight_decay=1.0e-5)
    b = FusedAdagrad(tst_p, lr=5e-4, eps=1e-8, weight_decay=1.0e-5)
    for _ in range(7):
        for r, t in zip(ref_p, tst_p):
            r.grad = torch.rand_like(r)
            t.grad = r.grad.clone()
        a.step(); b.step()
        assert max(float((r - t).abs().max()) for r, t in zip(ref_p, tst_p)) <= 1e-6
    assert "sum" in b.state[tst_p[0]]


def test_fused_novograd_vs_plain_pytorch():
    from apex.optimizers import FusedNovoGrad
    torch.manual_seed(3)
    sizes = [[4096, 64], [333], [1]]
    ref_p = [torch.rand(s, device="cuda", dtype=torch.float64) for s in sizes]
    tst_p = [torch.nn.Parameter(p.float()) for p in ref_p]
    ref_p = [p.float().double() for p in ref_p]
    lr, b1, b2, eps, wd = 1e-2, 0.95, 0.98, 1e-8, 0.01
    opt = FusedNovoGrad(tst_p, lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd, norm_type=2, reg_inside_moment=False)
    m = [torch.zeros_like(p) for p in ref_p]
    v = None
    for step in range(1, 5):
        grads = [torch.rand_like(p) for p in tst_p]
        for p, g in zip(tst_p, grads):
            p.grad = g.clone()
        opt.step()
        gn = torch.stack([g.double().norm() for g in grads])
        v = gn.clone() if v is None else torch.sqrt(b2 * v * v + (1 - b2) * gn * gn)   # first blend is a no-op
        if step == 1:
            v = torch.sqrt(b2 * gn * gn + (1 - b2) * gn * gn)
        bc1, bc2 = 1 - b1 ** step, (1 - b2 ** step) ** 0.5
        for i, (p, g) in enumerate(zip(ref_p, grads)):
            m[i] = b1 * m[i] + (1 - b1) * g.double()
            p -= lr * ((m[i] / bc1) / (v[i] / bc2 + eps) + wd * p)
        assert max(float((r.float() - t.detach()).abs().max()) for r, t in zip(ref_p, tst_p)) <= 1e-5
    assert opt.param_groups[0]["exp_avg_sq"][1].numel() == 3 and opt.param_groups[0]["exp_avg_sq"][0].numel() == 0


def test_e5m2_pack_unpack_matches_reference():
    """maybe_cast_mt between float/half and e5m2 bytes (fused_adam_cuda_kernel.cu:395-463), bit-exact."""
    import fused_adam_cuda as ours
    refm = load_ref("fused_adam_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/fused_adam_cuda_ref.so not built")
    g = torch.Generator(device="cuda").manual_seed(60)
    sizes = [1, 7, 555, 65537, 33333]
    for src_dt in (torch.float32, torch.float16):
        src = [(torch.randn(n, generator=g, device="cuda") * 10.0 ** torch.randint(-3, 4, (1,), device="cuda").float()).to(src_dt) for n in sizes]
        src[2][5] = 0.0
        src[3][100] = 70000.0 if src_dt == torch.float32 else 60000.0     # beyond / near the fp16 range
        packed = [torch.zeros(n, dtype=torch.uint8, device="cuda") for n in sizes]
        rpacked = [torch.zeros_like(t) for t in packed]
        ours.maybe_cast_mt(CHUNK, flag_tensor(0), [src, packed])
        refm.maybe_cast_mt(CHUNK, flag_tensor(0), [src, rpacked])
        for a, b in zip(packed, rpacked):
            assert torch.equal(a, b)
        for out_dt in (torch.float32, torch.float16):
            un = [torch.zeros(n, dtype=out_dt, device="cuda") for n in sizes]
            run = [torch.zeros_like(t) for t in un]
            ours.maybe_cast_mt(CHUNK, flag_tensor(0), [packed, un])
            refm.maybe_cast_mt(CHUNK, flag_tensor(0), [rpacked, run])
            for a, b in zip(un, run):
                assert torch.equal(a.isnan(), b.isnan()) and torch.equal(a.nan_to_num(), b.nan_to_num())
    # byte tensors through the strided finite check: 0x7c = +inf in e5m2
    y = torch.full((10000,), 0x3c, dtype=torch.uint8, device="cuda")
    y[4096] = 0x7c
    for stride, expect in ((64, 1), (4097, 0)):
        f, rf = flag_tensor(0), flag_tensor(0)
        ours.strided_check_finite(f, y, stride, 1)
        refm.strided_check_finite(rf, y, stride, 1)
        assert int(f) == int(rf) == expect


def test_oracle_e5m2_is_pinned_by_the_reference_kernel():
    """oracle.pack_e5m2 / unpack_e5m2 against the reference's own convert<> (fused_adam_cuda_kernel.cu:395-463) run
    through its maybe_cast_mt: every one of the 65536 Half bit patterns, all 256 bytes, and fp32 values around
    every rounding boundary."""
    refm = load_ref("fused_adam_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/fused_adam_cuda_ref.so not built")
    halves = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.float16).cuda()
    out = torch.zeros(65536, dtype=torch.uint8, device="cuda")
    refm.maybe_cast_mt(CHUNK, flag_tensor(0), [[halves], [out]])
    want = O.pack_e5m2(halves.float().cpu().numpy())
    assert np.array_equal(out.cpu().numpy(), want)
    g = torch.Generator(device="cuda").manual_seed(61)
    x = torch.randn(1 << 18, generator=g, device="cuda") * torch.exp2(torch.randint(-20, 18, (1 << 18,), generator=g, device="cuda").float())
    x[:8] = torch.tensor([0.0, -0.0, float("inf"), -float("inf"), float("nan"), 65504.0, 61440.0, 6e-8], device="cuda")
    out = torch.zeros(x.numel(), dtype=torch.uint8, device="cuda")
    refm.maybe_cast_mt(CHUNK, flag_tensor(0), [[x], [out]])
    assert np.array_equal(out.cpu().numpy(), O.pack_e5m2(x.cpu().numpy()))
    b = torch.arange(256, dtype=torch.int32).to(torch.uint8).cuda()
    f = torch.zeros(256, device="cuda")
    refm.maybe_cast_mt(CHUNK, flag_tensor(0), [[b], [f]])
    got, want = f.cpu().numpy(), O.unpack_e5m2(b.cpu().numpy())
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(np.nan_to_num(got), np.nan_to_num(want))


@pytest.mark.parametrize("gdt", ["f16", "bf16"])
def test_dist_adam_e5m2_copy(gdt):
    """distributed_adam_cuda.multi_tensor_fused_adam with an e5m2 p_copy (the payload of e5m2_allgather=True,
    distributed_fused_adam.py:198,206): p/m/v bit-identical to the run with a 16-bit copy, bytes bit-identical
    to the oracle's packing of the fp32 result."""
    import distributed_adam_cuda as ours
    T = len(SIZES)
    lists = rand_lists(SIZES, [torch.float32, torch.float32, torch.float32, DT[gdt]], 62, positive=(2,))
    plain = [[t.clone() for t in l] for l in lists] + [[torch.zeros(n, dtype=DT[gdt], device="cuda") for n in SIZES]]
    packed = [[t.clone() for t in l] for l in lists] + [[torch.zeros(n, dtype=torch.uint8, device="cuda") for n in SIZES]]
    dev = "cuda"
    b1, b2 = torch.full((T,), 0.9, device=dev), torch.full((T,), 0.999, device=dev)
    eps, wd = torch.full((T,), 1e-8, device=dev), torch.linspace(0, 0.1, T, device=dev)
    bc = torch.ones(T, dtype=torch.int32, device=dev)
    for step in (1, 2):
        ours.multi_tensor_fused_adam(CHUNK, flag_tensor(), plain, b1, b2, bc, eps, wd, 1e-2, 2.0, step, 1)
        ours.multi_tensor_fused_adam(CHUNK, flag_tensor(), packed, b1, b2, bc, eps, wd, 1e-2, 2.0, step, 1)
    for d in range(3):
        assert all(torch.equal(a, b) for a, b in zip(plain[d], packed[d]))
    for p, c in zip(packed[0], packed[4]):
        assert np.array_equal(c.cpu().numpy(), O.pack_e5m2(p.cpu().numpy()))
    # misaligned views (head peel + scalar tail) give the same bytes
    base = [torch.zeros(n + 3, dtype=torch.uint8, device="cuda") for n in SIZES]
    views = [b[3:] for b in base]
    again = [[t.clone() for t in l] for l in lists] + [views]
    for step in (1, 2):
        ours.multi_tensor_fused_adam(CHUNK, flag_tensor(), again, b1, b2, bc, eps, wd, 1e-2, 2.0, step, 1)
    assert all(torch.equal(a, b) for a, b in zip(views, packed[4]))
