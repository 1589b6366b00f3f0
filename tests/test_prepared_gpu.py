"""GPU: prepared list sets (b200mt_lists_prepare / amp_C.prepare) and the launch trace (b200mt_trace_begin / _end).

A prepared handle must give exactly the results of the plain call -- it only skips the per-call host work (list walk,
table image, compare) -- for every kind of op it is used with (element-wise and norm-type tables are created on
first use), on several streams, inside a CUDA graph after a warm-up, and must refuse list sets it was not made for.
"""
import ctypes
import time

import pytest
import torch

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32


def _lists(T=300, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    sizes = [1 + (37 * i) % 5000 for i in range(T)]
    mk = lambda s=1.0, pos=False: [((torch.randn(n, generator=g, device="cuda") * s).abs() if pos else torch.randn(n, generator=g, device="cuda") * s)
                                   for n in sizes]
    return mk(), mk(0.02), mk(0.01), mk(0.001, True)


def _flag():
    return torch.zeros(1, dtype=torch.int32, device="cuda")


def test_prepared_matches_plain_calls_for_every_op_kind():
    import amp_C
    import _b200mt as mt
    g, p, m, v = _lists()
    plain = [[t.clone() for t in l] for l in (g, p, m, v)]
    f = _flag()
    h4 = amp_C.prepare([g, p, m, v])
    h2 = amp_C.prepare([g, g])
    h1 = amp_C.prepare([g])
    up0 = mt.lib.b200mt_table_uploads()
    for step in (1, 2, 3):
        amp_C.multi_tensor_scale(CHUNK, f, h2, 0.5)
        amp_C.multi_tensor_scale(CHUNK, f, [plain[0], plain[0]], 0.5)
        n1, pt1 = amp_C.multi_tensor_l2norm(CHUNK, f, h1, True)
        n2, pt2 = amp_C.multi_tensor_l2norm(CHUNK, f, [plain[0]], True)
        assert torch.equal(n1, n2) and torch.equal(pt1, pt2)
        n3, _ = amp_C.multi_tensor_l2norm(333, f, h1, False)         # another chunk size: another table of the same handle
        n4, _ = amp_C.multi_tensor_l2norm(333, f, [plain[0]], False)
        assert torch.equal(n3, n4)
        args = (1e-3, 0.9, 0.999, 1e-8, step, 1, 1, 0.01)
        amp_C.multi_tensor_adam(CHUNK, f, h4, *args)
        amp_C.multi_tensor_adam(CHUNK, f, plain, *args)
        lamb_args = (1e-2, 0.9, 0.999, 1e-6, step, 1, 0.01, 1, 1)
        amp_C.multi_tensor_lamb(CHUNK, f, h4, *lamb_args, n1, 1.0, False)      # LAMB's norm-type table, third kind for h4
        amp_C.multi_tensor_lamb(CHUNK, f, plain, *lamb_args, n2, 1.0, False)
        for a, b in zip((g, p, m, v), plain):
            assert all(torch.equal(x, y) for x, y in zip(a, b)), step
    assert int(f) == 0
    # one upload per (handle, kind of op) for the prepared calls + the plain calls' own cache entries, none per step
    assert mt.lib.b200mt_table_uploads() - up0 <= 5 + 5
    for h in (h1, h2, h4):
        h.release()
        h.release()           # idempotent


def test_prepared_handle_rejects_other_list_sets_and_dead_handles():
    import _b200mt as mt
    g, p, m, v = _lists(T=10)
    h = mt.prepare([g, p])
    other = mt.pack([g[:5], p[:5]], 2)
    other.struct.prepared = h.handle
    f = _flag()
    rc = mt.lib.b200mt_scale(other.ref(), CHUNK, f.data_ptr(), 1.0, mt.stream_ptr(0))
    assert rc != 0 and "made for another list set" in mt.lib.b200mt_last_error().decode()
    with pytest.raises(RuntimeError, match="depth"):
        import amp_C
        amp_C.multi_tensor_adam(CHUNK, f, h, 1e-3, 0.9, 0.999, 1e-8, 1, 1, 1, 0.0)     # 2 lists where 4 are needed
    handle = h.handle
    h.release()
    stale = mt.pack([g, p], 2)
    stale.struct.prepared = handle
    # a released handle is freed memory: the library can only catch it while the magic word is still readable, so this
    # is not asserted -- but a NULL-handle call on the same lists must still work
    stale.struct.prepared = None
    mt.check(mt.lib.b200mt_scale(stale.ref(), CHUNK, f.data_ptr(), 1.0, mt.stream_ptr(0)))


def test_prepared_on_two_streams_and_in_a_graph():
    import amp_C
    g, p, m, v = _lists(T=64, seed=3)
    want = [t * 0.25 for t in g]
    h = amp_C.prepare([g, g])
    f = _flag()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    s1.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        amp_C.multi_tensor_scale(CHUNK, f, h, 0.5)       # uploads the table on s1
    s2.wait_stream(s1)
    with torch.cuda.stream(s2):
        amp_C.multi_tensor_scale(CHUNK, f, h, 0.5)       # first use on s2: ordered after the upload
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(g, want))
    # captured after the warm-up above: no allocation, no upload inside the capture
    gs = torch.cuda.Stream()
    gs.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=gs):
        amp_C.multi_tensor_scale(CHUNK, f, h, 2.0)
    for _ in range(2):
        graph.replay()
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(g, [t * 4.0 for t in want]))
    h.release()


def test_prepared_removes_the_per_call_host_work_on_10000_tensors():
    """10,000 x 1,024-element tensors (BASELINE.json configs[4], small end): the plain call walks 20,000 tensors and
    builds and compares a 240 KB image per call; the prepared call must cost a small fraction of that on the host."""
    import amp_C
    big = torch.randn(10000 * 1024, device="cuda")
    out = torch.empty_like(big)
    xs, os_ = list(big.view(10000, 1024).unbind(0)), list(out.view(10000, 1024).unbind(0))
    f = _flag()
    h = amp_C.prepare([xs, os_])

    def host_us(call, reps=30):
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            call()
        t1 = time.perf_counter()          # host time to ISSUE the calls (the queue is short: ~15 us of GPU work each)
        torch.cuda.synchronize()
        return (t1 - t0) / reps * 1e6

    plain = host_us(lambda: amp_C.multi_tensor_scale(CHUNK, f, [xs, os_], 0.5))
    prep = host_us(lambda: amp_C.multi_tensor_scale(CHUNK, f, h, 0.5))
    assert torch.equal(out, big * 0.5)
    print("10k x 1K scale: plain %.1f us/call, prepared %.1f us/call" % (plain, prep))
    assert prep < 0.35 * plain and prep < 100.0, (plain, prep)
    h.release()


def test_launch_trace_names_and_times_every_kernel():
    import amp_C
    import _b200mt as mt
    g, p, m, v = _lists(T=50, seed=5)
    f = _flag()
    amp_C.multi_tensor_l2norm(CHUNK, f, [g], False)     # warm-up outside the trace
    with mt.trace(16) as tr:
        amp_C.multi_tensor_scale(CHUNK, f, [g, g], 1.0)
        amp_C.multi_tensor_l2norm(CHUNK, f, [g], False)
        amp_C.multi_tensor_adam(CHUNK, f, [g, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 1, 1, 1, 0.0)
    names = [n for n, _ in tr.launches]
    assert len(names) == 4, names      # the norm is two kernels (B200MT_NORM_FOLD=1 would make it one)
    assert "ScaleOp" in names[0] and "L2NormOp" in names[1] and "norm_cleanup" in names[2] and "AdamOp" in names[3]
    assert all(0.0 < ms < 5.0 for _, ms in tr.launches), tr.launches
    # tracing is off again: nothing is recorded, results unchanged
    before = mt.lib.b200mt_launch_count()
    amp_C.multi_tensor_scale(CHUNK, f, [g, g], 1.0)
    assert mt.lib.b200mt_launch_count() == before + 1
    rc = mt.lib.b200mt_trace_end((ctypes.c_float * 1)(), 1, None)
    assert rc != 0 and "no trace is running" in mt.lib.b200mt_last_error().decode()
