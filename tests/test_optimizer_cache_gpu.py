"""GPU: the memoised tensor lists of FusedAdam / FusedLAMB (apex/optimizers/_common.py: GroupCache) must notice everything
the reference's per-step Python loop (apex/optimizers/fused_adam.py:118-160) would have seen: gradients appearing and
disappearing, state loaded from a checkpoint, state tensors replaced by the user, parameter groups added, gradient
tensors re-allocated every step.  FusedAdam is stepped next to torch.optim.AdamW in the steady state and next to an uncached twin through all of it."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def make(n=6, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    shapes = [(33, 7), (128,), (5,), (64, 64), (1,), (1000,)][:n]
    a = [torch.nn.Parameter(torch.randn(s, generator=g, device="cuda") * 0.1) for s in shapes]
    b = [torch.nn.Parameter(p.detach().clone()) for p in a]
    return a, b


def set_grads(a, b, seed, skip=()):
    g = torch.Generator(device="cuda").manual_seed(seed)
    for i, (p, q) in enumerate(zip(a, b)):
        gr = torch.randn(p.shape, generator=g, device="cuda")
        if i in skip:
            p.grad, q.grad = None, None
        else:
            p.grad, q.grad = gr.clone(), gr.clone()      # fresh tensors every step, like backward after zero_grad(set_to_none)


def same(a, b, tol=2e-6):
    return all(float((p.detach() - q.detach()).abs().max()) <= tol for p, q in zip(a, b))


def test_fused_adam_cache_tracks_grads_state_and_groups():
    from apex.optimizers import FusedAdam
    a, b = make()
    kw = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01)
    # steady state against torch.optim.AdamW first (same per-step counters as long as every parameter has a gradient)
    x, y = make(seed=1)
    fx, ty = FusedAdam(x, **kw), torch.optim.AdamW(y, **kw)
    for it in range(4):
        set_grads(x, y, it)
        fx.step(); ty.step()
        assert same(x, y)

    class Uncached(FusedAdam):                            # the reference's loop on every step
        def step(self, *args, **kwargs):
            self._lists.clear()
            return super().step(*args, **kwargs)

    fa, ta = FusedAdam(a[:4], **kw), Uncached(b[:4], **kw)
    hits = lambda: fa._lists.groups[0].buckets
    for it in range(3):                                   # cache built on step 0, reused after
        set_grads(a, b, it)
        fa.step(); ta.step()
        assert same(a, b, tol=0.0)
    rows_before = hits()
    set_grads(a, b, 10)
    fa.step(); ta.step()
    assert hits() is rows_before and same(a, b)           # fast path really was taken
    # a parameter without gradient this step, then with one again
    set_grads(a, b, 11, skip=(1,))
    fa.step(); ta.step()
    assert same(a, b) and hits() is not rows_before
    set_grads(a, b, 12)
    fa.step(); ta.step()
    assert same(a, b)
    # a new parameter group
    fa.add_param_group({"params": a[4:], "lr": 5e-3})
    ta.add_param_group({"params": b[4:], "lr": 5e-3})
    for it in range(2):
        set_grads(a, b, 20 + it)
        fa.step(); ta.step()
        assert same(a, b)
    # the user replaces a state tensor behind the optimizer's back
    fa.state[a[0]]["exp_avg"] = torch.zeros_like(a[0])
    ta.state[b[0]]["exp_avg"] = torch.zeros_like(b[0])
    set_grads(a, b, 30)
    fa.step(); ta.step()
    assert same(a, b)
    # checkpoint round trip into a fresh optimizer on cloned parameters, both continue identically
    a2 = [torch.nn.Parameter(p.detach().clone()) for p in a]
    fa2 = FusedAdam([{"params": a2[:4]}, {"params": a2[4:], "lr": 5e-3}], **kw)
    set_grads(a2, a2, 0)
    fa2.step()                                            # builds a cache that the load below must invalidate
    with torch.no_grad():
        for p, q in zip(a2, a):
            p.copy_(q)
    fa2.load_state_dict(copy.deepcopy(fa.state_dict()))
    for it in range(2):
        set_grads(a, b, 40 + it)
        for p, q in zip(a2, a):
            p.grad = q.grad.clone()
        fa.step(); ta.step(); fa2.step()
        assert same(a, b) and same(a2, a, tol=0.0)
    # zero_grad(set_to_none) then a step with no gradients at all is a no-op
    fa.zero_grad()
    before = [p.detach().clone() for p in a]
    fa.step()
    assert all(torch.equal(x, p.detach()) for x, p in zip(before, a))


def test_fused_lamb_cache_matches_uncached_steps():
    from apex.optimizers import FusedLAMB
    a, b = make()
    kw = dict(lr=1e-2, weight_decay=0.01, max_grad_norm=1.0)
    cached, fresh = FusedLAMB(a, **kw), FusedLAMB(b, **kw)
    for it in range(5):
        set_grads(a, b, it, skip=(2,) if it == 3 else ())
        fresh._lists.clear()                              # always the slow pass: the reference's loop
        cached.step(); fresh.step()
        assert same(a, b, tol=0.0)
    assert cached.param_groups[0]["step"] == 5


def test_fused_sgd_memo_matches_uncached_steps_and_torch():
    from apex.optimizers import FusedSGD
    a, b = make()
    c = [torch.nn.Parameter(p.detach().clone()) for p in a]
    kw = dict(lr=1e-2, momentum=0.9, weight_decay=1e-4, nesterov=True)
    cached, fresh, ref = FusedSGD(a, **kw), FusedSGD(b, **kw), torch.optim.SGD(c, **kw)
    for it in range(6):
        skip = (2,) if it == 3 else ()
        set_grads(a, b, it, skip=skip)
        for p, q in zip(c, a):
            p.grad = None if q.grad is None else q.grad.clone()
        fresh._sets.clear()                               # always the reference's three walks over the parameters
        cached.step(); fresh.step(); ref.step()
        assert same(a, b, tol=0.0) and same(a, c, tol=1e-6)
        if it == 2:
            assert 0 in cached._sets                      # memo in place once the momentum buffers exist
    sd = copy.deepcopy(cached.state_dict())
    cached.load_state_dict(sd)                            # state objects replaced: the memo must not survive
    assert cached._sets == {}
    set_grads(a, b, 50)
    cached.step(); fresh._sets.clear(); fresh.step()
    assert same(a, b, tol=0.0)


def test_fused_adam_prepared_handle_follows_the_gradients():
    """FusedAdam hands its kernel a prepared list handle while the gradients stay where they were (one C loop checks the
    addresses and the state tensors), re-prepares when they move, gives up preparing a bucket that moves every step, and
    never uses a stale address: stepped next to torch.optim.AdamW through all of it."""
    import _b200mt as mt
    from apex.optimizers import FusedAdam
    if mt.check_rows is None:
        pytest.skip("native packer not built")
    kw = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01)
    x, y = make(seed=3)
    fx, ty = FusedAdam(x, **kw), torch.optim.AdamW(y, **kw)
    handle = lambda: fx._lists.groups[0].buckets[0].handles.get("prep")
    g = torch.Generator(device="cuda").manual_seed(5)
    for p, q in zip(x, y):
        p.grad = torch.randn(p.shape, generator=g, device="cuda")
        q.grad = p.grad.clone()
    fx.step(); ty.step()
    first = handle()
    assert first is not None and first.keep is None
    up0, hit0 = mt.lib.b200mt_table_uploads(), mt.lib.b200mt_table_hits()
    for it in range(4):                                   # gradients rewritten in place: same addresses, same handle
        for p, q in zip(x, y):
            p.grad.copy_(torch.randn(p.shape, generator=g, device="cuda"))
            q.grad.copy_(p.grad)
        fx.step(); ty.step()
        assert same(x, y, tol=1e-5) and handle() is first
    assert mt.lib.b200mt_table_uploads() == up0 and mt.lib.b200mt_table_hits() == hit0 + 4
    # one gradient moves: the handle is rebuilt, results still right
    x[2].grad = torch.randn(x[2].shape, generator=g, device="cuda"); y[2].grad = x[2].grad.clone()
    fx.step(); ty.step()
    assert same(x, y, tol=1e-5) and handle() is not first and handle() is not None
    # a state tensor replaced by the user: noticed (rc 2), the new tensor is used
    fx.state[x[1]]["exp_avg"] = fx.state[x[1]]["exp_avg"].clone()
    ty_state = ty.state[y[1]]["exp_avg"]
    fx.step(); ty.step()
    assert same(x, y, tol=1e-5) and ty_state is ty.state[y[1]]["exp_avg"]
    # gradients that land somewhere else every step: after three misses the bucket goes back to plain lists
    keep = []
    for it in range(5):
        for p, q in zip(x, y):
            p.grad = torch.randn(p.shape, generator=g, device="cuda"); q.grad = p.grad.clone()
            keep.append(p.grad)                           # held, so the allocator cannot hand the address out again
        fx.step(); ty.step()
        assert same(x, y, tol=1e-5)
    assert handle() is None
    # a parameter without gradient this step, then with one again
    x[0].grad, y[0].grad = None, None
    fx.step(); ty.step()
    for p, q in zip(x, y):
        p.grad = torch.randn(p.shape, generator=g, device="cuda"); q.grad = p.grad.clone()
    fx.step(); ty.step()
    assert same(x[1:], y[1:], tol=1e-5)


def test_fused_sgd_prepared_handle_follows_the_gradients():
    """FusedSGD's memoised [g, p, momentum] sets go to the kernel as a prepared handle while the gradients stay put, are
    re-prepared when they move and fall back to plain lists when they keep moving; next to torch.optim.SGD throughout."""
    import _b200mt as mt
    from apex.optimizers import FusedSGD
    if mt.check_rows is None:
        pytest.skip("native packer not built")
    kw = dict(lr=0.05, momentum=0.9, weight_decay=0.01, nesterov=True)
    x, y = make(seed=4)
    fx, ty = FusedSGD(x, **kw), torch.optim.SGD(y, **kw)
    g = torch.Generator(device="cuda").manual_seed(6)
    for p, q in zip(x, y):
        p.grad = torch.randn(p.shape, generator=g, device="cuda"); q.grad = p.grad.clone()
    for it in range(2):                                   # step 0 creates the momentum buffers, step 1 builds the memo
        fx.step(); ty.step()
    handle = lambda: fx._sets[0].buckets[1].handles.get("prep")    # the fp32 bucket
    hits0 = None
    for it in range(4):
        for p, q in zip(x, y):
            p.grad.copy_(torch.randn(p.shape, generator=g, device="cuda")); q.grad.copy_(p.grad)
        fx.step(); ty.step()
        assert same(x, y, tol=1e-5)
        assert handle() is not None
        if hits0 is None:
            first, hits0 = handle(), mt.lib.b200mt_table_hits()
    assert handle() is first and mt.lib.b200mt_table_hits() >= hits0 + 3
    keep = []
    for it in range(5):                                   # gradients somewhere else every step
        for p, q in zip(x, y):
            p.grad = torch.randn(p.shape, generator=g, device="cuda"); q.grad = p.grad.clone()
            keep.append(p.grad)
        fx.step(); ty.step()
        assert same(x, y, tol=1e-5)
    assert handle() is None
    fx.state[x[3]]["momentum_buffer"] = fx.state[x[3]]["momentum_buffer"].clone()      # state replaced: memo rebuilt
    fx.step(); ty.step()
    assert same(x, y, tol=1e-5)


@pytest.mark.parametrize("two_groups", [False, True])
def test_fused_lamb_prepared_handles_match_plain_lists(two_groups):
    """FusedLAMB norms the gradients through a depth-1 handle prepared together with the [g, p, m, v] one (a dtype that
    sits in one bucket) or as one concatenated list (a dtype spread over two groups); a twin that is only ever given
    plain lists must end up bit-identical, through gradients rewritten in place, moved, and missing."""
    import _b200mt as mt
    from apex.optimizers import FusedLAMB
    if mt.check_rows is None:
        pytest.skip("native packer not built")
    x, y = make(seed=7)
    h16 = torch.nn.Parameter((torch.randn(300, device="cuda") * 0.1).half())
    x, y = x + [h16], y + [torch.nn.Parameter(h16.detach().clone())]
    kw = dict(lr=1e-2, weight_decay=0.01, max_grad_norm=0.5)
    groups = lambda ps: [{"params": ps[:4] + ps[6:]}, {"params": ps[4:6], "weight_decay": 0.0}] if two_groups else ps
    fx, fy = FusedLAMB(groups(x), **kw), FusedLAMB(groups(y), **kw)
    fy._lists._maybe_prepare = lambda bucket, gs: [gs, bucket.params, bucket.state0, bucket.state1]     # the twin never prepares
    g = torch.Generator(device="cuda").manual_seed(8)

    def fresh_grads(keep=None):
        for p, q in zip(x, y):
            p.grad = torch.randn(p.shape, generator=g, device="cuda").to(p.dtype); q.grad = p.grad.clone()
            if keep is not None:
                keep.append(p.grad)

    fresh_grads()
    fx.step(); fy.step()
    assert same(x, y, tol=0.0)
    b0 = fx._lists.groups[0].buckets
    assert [b.params[0].dtype for b in b0] == [torch.float32, torch.float16]
    assert all(b.handles.get("prep") is not None and b.handles.get("prep_g") is not None and b.handles["prep_g"].depth == 1
               for b in b0)
    assert all(b.handles.get("prep") is None for b in fy._lists.groups[0].buckets)
    first = [b.handles["prep_g"] for b in b0]
    up0 = mt.lib.b200mt_table_uploads()
    for it in range(3):                                   # in place (the update kernel has overwritten g): same handles
        for p, q in zip(x, y):
            p.grad.copy_(torch.randn(p.shape, generator=g, device="cuda")); q.grad.copy_(p.grad)
        fx.step(); fy.step()
        assert same(x, y, tol=0.0)
        assert [b.handles["prep_g"] for b in fx._lists.groups[0].buckets] == first
    if not two_groups:                                    # nothing but the list of two norms can have been uploaded
        assert mt.lib.b200mt_table_uploads() - up0 <= 3
    keep = []
    for it in range(5):                                   # gradients somewhere else every step: back to plain lists
        fresh_grads(keep)
        fx.step(); fy.step()
        assert same(x, y, tol=0.0)
    assert all(b.handles.get("prep") is None and b.handles.get("prep_g") is None for b in fx._lists.groups[0].buckets)
    x[1].grad, y[1].grad = None, None                     # a parameter without gradient, then with one again
    fx.step(); fy.step()
    fresh_grads()
    fx.step(); fy.step()
    assert same(x, y, tol=0.0)
    fx._lists.clear()
