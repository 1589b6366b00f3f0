"""GPU: every op runs asynchronously on torch's CURRENT stream of the tensors' device (the reference launches on
at::cuda::getCurrentCUDAStream(), csrc/multi_tensor_apply.cuh:78-79; apex DDP calls multi_tensor_scale inside
torch.cuda.stream(bucket_stream), apex/parallel/distributed.py:437-465).  The cached descriptor table of a list
must stay coherent when the same list is used from several streams, and when the cache wraps around."""
import pytest
import torch

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32


def flag_tensor(v=0):
    return torch.full((1,), v, dtype=torch.int32, device="cuda")


def test_ops_follow_the_current_stream_and_share_tables_across_streams():
    import amp_C
    import _b200mt as mt
    n = 1 << 22
    xs = [torch.zeros(n, device="cuda") for _ in range(4)]
    f = flag_tensor()
    streams = [torch.cuda.Stream() for _ in range(3)]
    torch.cuda.synchronize()
    expected = 0.0
    for it in range(12):
        s = streams[it % 3]
        prev = streams[(it - 1) % 3]
        s.wait_stream(prev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            # a slow producer on this stream, then our op on the same list (same cached table, different stream each time)
            for x in xs:
                x.add_(1.0)
            amp_C.multi_tensor_scale(CHUNK, f, [xs, xs], 2.0)
        expected = (expected + 1.0) * 2.0
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    assert all(float(x.min()) == float(x.max()) == expected for x in xs)
    assert mt.lib.b200mt_table_hits() > 0


def test_table_cache_wraps_without_corruption():
    import amp_C
    # far more distinct list sets than cache slots (64), interleaved with a re-used one
    f = flag_tensor()
    base = [torch.full((5000 + 17 * i,), float(i), device="cuda") for i in range(200)]
    keep = [torch.ones(70000, device="cuda"), torch.ones(3, device="cuda")]
    for rnd in range(2):
        for i, t in enumerate(base):
            out = torch.empty_like(t)
            amp_C.multi_tensor_scale(CHUNK, f, [[t, keep[1]], [out, torch.empty_like(keep[1])]], 3.0)
            if i % 7 == 0:
                amp_C.multi_tensor_scale(CHUNK, f, [keep, keep], 1.0)
            assert float(out[0]) == 3.0 * i and float(out[-1]) == 3.0 * i
    n, per = amp_C.multi_tensor_l2norm(CHUNK, f, [keep], True)
    assert abs(float(n) - (70003 ** 0.5)) < 1e-3 and int(f) == 0


def test_l2norm_scratch_is_per_stream():
    import amp_C
    xs = [torch.full((1 << 20,), 2.0, device="cuda") for _ in range(8)]
    ys = [torch.full((1 << 20,), 3.0, device="cuda") for _ in range(8)]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    res = []
    for _ in range(5):
        with torch.cuda.stream(s1):
            a = amp_C.multi_tensor_l2norm(CHUNK, flag_tensor(), [xs], True)
        with torch.cuda.stream(s2):
            b = amp_C.multi_tensor_l2norm(CHUNK, flag_tensor(), [ys], True)
        res.append((a, b))
    torch.cuda.synchronize()
    for (an, ap), (bn, bp) in res:
        assert abs(float(an) - 2.0 * (8 << 20) ** 0.5) < 1e-2 and abs(float(bn) - 3.0 * (8 << 20) ** 0.5) < 1e-2
        assert torch.allclose(ap, torch.full_like(ap, 2.0 * 1024.0)) and torch.allclose(bp, torch.full_like(bp, 3.0 * 1024.0))


def test_host_threads_on_their_own_streams_share_the_library():
    """Four host threads, each on its own stream, hammer the same library instance (ctypes drops the GIL inside a call, so the
    calls really overlap): per-device lock, table cache (more distinct lists than it has slots, so slots are evicted under
    the other threads' feet), per-stream scratch of the norm ops.  Every thread must see exactly what it would have seen alone."""
    import threading

    import amp_C
    nthreads, rounds = 4, 60
    errors = []

    def worker(tid):
        try:
            torch.cuda.set_device(0)
            stream = torch.cuda.Stream()
            g = torch.Generator(device="cuda").manual_seed(100 + tid)
            with torch.cuda.stream(stream):
                f = torch.zeros(1, dtype=torch.int32, device="cuda")
                # 30 list sets per thread -> 120 in flight against 64 cache slots
                sets = []
                for k in range(30):
                    sizes = [1000 + 37 * k + 11 * tid, 5 + k, 70000 + tid]
                    xs = [torch.randn(n, generator=g, device="cuda") for n in sizes]
                    sets.append((xs, [torch.empty_like(x) for x in xs], float(torch.cat(xs).double().norm())))
                for r in range(rounds):
                    xs, outs, want = sets[(r * 7 + tid) % len(sets)]
                    s = 0.5 + tid
                    amp_C.multi_tensor_scale(CHUNK, f, [xs, outs], s)
                    n, per = amp_C.multi_tensor_l2norm(CHUNK, f, [xs], True)
                    n2, _ = amp_C.multi_tensor_l2norm(CHUNK, f, [outs], False)
                    got, got2 = float(n), float(n2)           # synchronises this thread's stream only
                    assert abs(got - want) <= 1e-5 * want, (tid, r, got, want)
                    assert abs(got2 - s * want) <= 1e-5 * s * want, (tid, r, got2, s * want)
                    for x, o, pn in zip(xs, outs, per.tolist()):
                        assert torch.equal(o, x * s)
                        assert abs(pn - float(x.double().norm())) <= 1e-5 * max(pn, 1e-30)
                assert int(f) == 0
        except Exception as e:                                # surfaces in the main thread
            errors.append((tid, repr(e)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(nthreads)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    torch.cuda.synchronize()
    assert not any(t.is_alive() for t in threads), "a worker hung"
    assert errors == [], errors
