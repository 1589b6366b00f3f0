"""GPU: the C-ABI ops inside a CUDA graph (torch.cuda.graph).  The reference's device-resident-state ops
(multi_tensor_l2norm_mp + multi_tensor_lamb_mp: step, lr, loss scale and overflow flag all live on the device,
csrc/multi_tensor_lamb_mp.cu:367-496) exist to be captured; the descriptor-table cache and the scratch allocator must
therefore neither allocate, synchronise nor cross the capture boundary while a stream is capturing, and the table a
graph points at must never be evicted.  Replays must give the bits of the eager run."""
import pytest
import torch

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32
SIZES = [1, 3, 555, 2048 * 32 + 1, 33333, 8192, 16, 2048 * 32 * 2 - 1, 70001, 1024, 1024, 5]


def make_state(seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda dt, s: [(torch.randn(n, generator=g, device="cuda") * s).to(dt) for n in SIZES]
    st = dict(g16=mk(torch.float16, 1.0), p=mk(torch.float32, 0.05), m=mk(torch.float32, 0.01),
              v=[t.abs() for t in mk(torch.float32, 0.001)])
    st["p16"] = [t.half() for t in st["p"]]
    one = lambda val, dt=torch.float32: torch.tensor([val], dtype=dt, device="cuda")
    st.update(lr=one(1e-2), step=one(1, torch.int32), inv_scale=one(0.5), found_inf=one(0.0), max_norm=one(1.0),
              flag=torch.zeros(1, dtype=torch.int32, device="cuda"))
    return st


def lamb_step(amp_C, st):
    gn, _ = amp_C.multi_tensor_l2norm_mp(CHUNK, st["flag"], [st["g16"]], False)
    amp_C.multi_tensor_lamb_mp(CHUNK, st["flag"], [st["g16"], st["p"], st["m"], st["v"], st["p16"]], st["lr"], 0.9, 0.999,
                               1e-6, st["step"], 1, 0.01, 1, 1, gn, st["max_norm"], False, st["found_inf"], st["inv_scale"])
    st["step"] += 1


def clone_state(st):
    return {k: ([t.clone() for t in v] if isinstance(v, list) else v.clone()) for k, v in st.items()}


def restore(dst, src):
    for k, v in src.items():
        if isinstance(v, list):
            for a, b in zip(dst[k], v):
                a.copy_(b)
        else:
            dst[k].copy_(v)


def test_lamb_mp_step_replays_bit_exactly_from_a_cuda_graph():
    import amp_C
    import _b200mt as mt
    init = make_state(5)
    eager = clone_state(init)
    grads = [[(g * (1.0 + 0.1 * it)).half() for g in init["g16"]] for it in range(3)]   # the update consumes g: refill per step
    for it in range(3):
        for a, b in zip(eager["g16"], grads[it]):
            a.copy_(b)
        lamb_step(amp_C, eager)

    st = clone_state(init)
    static_in = [t.clone() for t in grads[0]]
    # warm-up (executes step 1) on a side stream, as torch's CUDA-graph recipe does
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for a, b in zip(st["g16"], static_in):
            a.copy_(b)
        lamb_step(amp_C, st)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    uploads = mt.lib.b200mt_table_uploads()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):                 # torch's own capture stream: never seen by the library before
        for a, b in zip(st["g16"], static_in):
            a.copy_(b)
        lamb_step(amp_C, st)
    assert mt.lib.b200mt_table_uploads() == uploads      # nothing was uploaded while capturing
    for it in (1, 2):
        for a, b in zip(static_in, grads[it]):
            a.copy_(b)
        graph.replay()
    torch.cuda.synchronize()
    assert int(st["step"]) == int(eager["step"]) == 4
    for k in ("p", "m", "v", "p16"):
        for a, b in zip(st[k], eager[k]):
            assert torch.equal(a, b), k
    # the tables the graph points at survive any amount of cache churn (64 slots, LRU)
    for i in range(80):
        x = torch.ones(64 + i, device="cuda")
        amp_C.multi_tensor_scale(CHUNK, st["flag"], [[x], [x]], 1.0)
    restore(st, init)
    restore(eager, init)
    for a, b in zip(static_in, grads[0]):
        a.copy_(b)
    graph.replay()
    for a, b in zip(eager["g16"], grads[0]):
        a.copy_(b)
    lamb_step(amp_C, eager)
    torch.cuda.synchronize()
    for k in ("p", "m", "v", "p16"):
        for a, b in zip(st[k], eager[k]):
            assert torch.equal(a, b), k


def test_overflow_inside_a_replay_skips_the_step():
    import amp_C
    st = make_state(6)
    keep = clone_state(st)
    static_in = [t.clone() for t in st["g16"]]

    def body():
        st["flag"].zero_()
        for a, b in zip(st["g16"], static_in):
            a.copy_(b)
        amp_C.multi_tensor_scale(CHUNK, st["flag"], [st["g16"], st["g16"]], 1.0)     # amp's overflow check
        lamb_step(amp_C, st)

    body()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        body()
    restore(st, keep)
    static_in[3][17] = float("inf")
    graph.replay()
    torch.cuda.synchronize()
    assert int(st["flag"]) == 1
    for k in ("p", "m", "v", "p16"):           # every kernel behind the flag returned at once
        for a, b in zip(st[k], keep[k]):
            assert torch.equal(a, b), k


def test_unseen_list_during_capture_is_refused_not_uploaded():
    import amp_C
    x = [torch.randn(1000 + i, device="cuda") for i in range(7)]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        graph.capture_begin()
        try:
            with pytest.raises(RuntimeError, match="stream capture"):
                amp_C.multi_tensor_scale(CHUNK, f, [x, x], 2.0)
        finally:
            graph.capture_end()
    torch.cuda.current_stream().wait_stream(s)
    keep = [t.clone() for t in x]
    amp_C.multi_tensor_scale(CHUNK, f, [x, x], 2.0)          # the library is unharmed
    assert all(torch.equal(a, b * 2.0) for a, b in zip(x, keep))
