"""GPU: b200mt_l2norm_flat_noalloc, the entry point for hosts that cannot let the library allocate
(their GPU memory already belongs to another runtime's pool, or a cudaMalloc in the step loop would
stall the device).  Its working memory is a fixed array inside the library, shared by every call
and never cleared between them, so the test calls it twice in a row -- second time on different
data -- and checks each result against the regular b200mt_l2norm entry point through the C ABI.

Collected last on purpose (file name): a run that stops at the first failure has then already
exercised every other test.  It passes alone in a fresh process and after the other tests.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

CHUNK = 2048 * 32
RTOL = 1e-5   # the norm tolerance used throughout tests/test_parity_gpu.py


def _regular(mt, x, chunk, flag):
    ret = torch.empty(1, device=x.device)
    L = mt.pack([[x]], 1)
    mt.check(mt.lib.b200mt_l2norm(L.ref(), chunk, flag.data_ptr(), ret.data_ptr(), None, 0,
                                  mt.stream_ptr(L.device)))
    return ret


def _noalloc(mt, x, chunk, flag):
    ret = torch.empty(1, device=x.device)
    dev = x.device.index
    mt.check(mt.lib.b200mt_l2norm_flat_noalloc(x.data_ptr(), mt.dtype_code(x), x.numel(), chunk, flag.data_ptr(),
                                               ret.data_ptr(), dev, mt.stream_ptr(dev)))
    return ret


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
def test_noalloc_norm_matches_regular_entry_point_twice_in_a_row(dtype):
    import _b200mt as mt
    g = torch.Generator(device="cuda").manual_seed(1234)
    a = (torch.randn(3_000_001, generator=g, device="cuda") * 3.0).to(dtype)
    b = (torch.randn(70_003, generator=g, device="cuda") * 0.01).to(dtype)[3:]   # smaller, misaligned view
    for x in (a, b):          # two calls in a row: the second one finds the first one's leftovers
        f0 = torch.zeros(1, dtype=torch.int32, device="cuda")
        f1 = torch.zeros(1, dtype=torch.int32, device="cuda")
        want = _regular(mt, x, CHUNK, f0)
        got = _noalloc(mt, x, CHUNK, f1)
        torch.cuda.synchronize()
        assert int(f0) == int(f1) == 0
        assert abs(float(got) - float(want)) <= RTOL * abs(float(want)), (float(got), float(want))
        assert abs(float(got) - float(x.float().norm())) <= 1e-4 * float(x.float().norm())
        # and against the reference's own kernel (oracle/_ref/amp_C_ref.so), when it was built
        from util import load_ref
        refm = load_ref("amp_C_ref")
        if refm is not None:
            rn, _ = refm.multi_tensor_l2norm(CHUNK, torch.zeros(1, dtype=torch.int32, device="cuda"), [[x]], False)
            assert abs(float(got) - float(rn)) <= RTOL * abs(float(rn)), (float(got), float(rn))


def test_noalloc_flag_and_reuse_after_overflow():
    import _b200mt as mt
    x = torch.ones(200_000, device="cuda")
    x[123_456] = float("inf")
    f0 = torch.zeros(1, dtype=torch.int32, device="cuda")
    f1 = torch.zeros(1, dtype=torch.int32, device="cuda")
    _regular(mt, x, 512, f0)
    _noalloc(mt, x, 512, f1)
    assert int(f0) == int(f1) == 1
    # the shared array now holds an inf partial: a clean input right after must not see it
    y = torch.full((4096,), 2.0, device="cuda")
    f2 = torch.zeros(1, dtype=torch.int32, device="cuda")
    got = _noalloc(mt, y, 512, f2)
    assert int(f2) == 0 and abs(float(got) - 128.0) <= RTOL * 128.0
    # empty input: norm 0, no flag
    f3 = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert float(_noalloc(mt, y[:0], 512, f3)) == 0.0 and int(f3) == 0


def test_noalloc_rejects_inputs_beyond_the_fixed_scratch():
    import _b200mt as mt
    x = torch.ones(70_000, device="cuda")
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    ret = torch.empty(1, device="cuda")
    # chunk_size 1 -> one tile per element -> 70000 tiles > 65536
    rc = mt.lib.b200mt_l2norm_flat_noalloc(x.data_ptr(), mt.F32, x.numel(), 1, f.data_ptr(), ret.data_ptr(), 0,
                                           mt.stream_ptr(0))
    assert rc != 0
    msg = mt.lib.b200mt_last_error().decode()
    assert "fixed scratch holds 65536" in msg and "70000 tiles" in msg


def test_noalloc_limit_at_a_chunk_size_that_is_not_a_multiple_of_the_tile():
    """chunk_size 16400 is cut into 2 tiles of 8208: the limit is 32768 chunks = 537,395,200 elements, not the
    1,073,741,824 of chunk sizes that are multiples of 16384.  Exactly at the limit the call works; one element more
    is refused on the host."""
    import _b200mt as mt
    chunk = 16400
    n = 32768 * chunk
    x = torch.ones(n + 1, dtype=torch.float16, device="cuda")
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    got = _noalloc(mt, x[:n], chunk, f)
    want = float(n) ** 0.5
    assert int(f) == 0 and abs(float(got) - want) <= RTOL * want, (float(got), want)
    ret = torch.empty(1, device="cuda")
    rc = mt.lib.b200mt_l2norm_flat_noalloc(x.data_ptr(), mt.F16, n + 1, chunk, f.data_ptr(), ret.data_ptr(), 0, mt.stream_ptr(0))
    assert rc != 0
    msg = mt.lib.b200mt_last_error().decode()
    assert "65537 tiles" in msg and "tile 8208 elements" in msg, msg
