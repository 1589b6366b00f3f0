"""CPU: the list memo of the fused optimizers (apex/optimizers/_common.py: GroupCache) is host logic only -- no kernel is
involved in deciding whether the lists of the previous step may be reused -- so its invalidation rules are checked here
with plain CPU tensors and a stand-in optimizer object.  The rules replace the reference's per-step loop over every
parameter (apex/optimizers/fused_adam.py:118-160); anything the loop would have noticed must force the slow pass."""
import os
import sys
import types
from collections import defaultdict

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def GroupCache():
    # the package __init__ needs the CUDA library; the memo does not -- load the module file on its own
    import importlib.util
    pkg = types.ModuleType("apex_stub")
    pkg.__path__ = []
    sys.modules.setdefault("apex_stub", pkg)
    mta = types.ModuleType("apex_stub.multi_tensor_apply")
    mta.multi_tensor_applier = None
    sys.modules["apex_stub.multi_tensor_apply"] = mta
    opt_pkg = types.ModuleType("apex_stub.optimizers")
    opt_pkg.__path__ = []
    sys.modules["apex_stub.optimizers"] = opt_pkg
    spec = importlib.util.spec_from_file_location(
        "apex_stub.optimizers._common", os.path.join(ROOT, "rocm-apex_b200", "apex", "optimizers", "_common.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["apex_stub.optimizers._common"] = mod
    spec.loader.exec_module(mod)
    return mod.GroupCache


def make(n=5, dtypes=(torch.float32,)):
    params = [torch.nn.Parameter(torch.zeros(3 + i, dtype=dtypes[i % len(dtypes)])) for i in range(n)]
    for p in params:
        p.grad = torch.ones_like(p)
    opt = types.SimpleNamespace(state=defaultdict(dict))
    group = {"params": params}
    return opt, group, params


def init_state(p, st):
    st["exp_avg"] = torch.zeros_like(p)
    st["exp_avg_sq"] = torch.zeros_like(p)


def test_steady_state_reuses_lists_and_rereads_grads(GroupCache):
    c = GroupCache("FusedAdam", ("exp_avg", "exp_avg_sq"))
    opt, group, params = make()
    b1 = c.buckets(opt, 0, group, init_state)
    assert len(b1) == 1 and [len(col) for col in b1[0]] == [5, 5, 5, 5]
    assert all(a is p.grad for a, p in zip(b1[0][0], params)) and all(a is p for a, p in zip(b1[0][1], params))
    assert all(a is opt.state[p]["exp_avg"] for a, p in zip(b1[0][2], params))
    for p in params:                       # backward after zero_grad(set_to_none=True): new gradient tensors
        p.grad = torch.full_like(p, 2.0)
    rows = c.groups[0].buckets
    b2 = c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is rows                                   # memo kept
    assert all(a is p.grad for a, p in zip(b2[0][0], params))       # gradients are the new ones
    assert b2[0][1] is b1[0][1] and b2[0][2] is b1[0][2]            # the other columns are the memoised lists


def test_every_change_the_reference_loop_would_see_forces_the_slow_pass(GroupCache):
    c = GroupCache("FusedAdam", ("exp_avg", "exp_avg_sq"))
    opt, group, params = make()
    c.buckets(opt, 0, group, init_state)
    rows = c.groups[0].buckets

    params[2].grad = None                                           # a gradient disappears
    b = c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is not rows and len(b[0][0]) == 4 and all(q is not params[2] for q in b[0][1])
    rows = c.groups[0].buckets
    params[2].grad = torch.ones_like(params[2])                     # ... and comes back
    b = c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is not rows and len(b[0][0]) == 5
    rows = c.groups[0].buckets

    new_m = torch.zeros_like(params[0])                             # a state tensor is replaced in place
    opt.state[params[0]]["exp_avg"] = new_m
    b = c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is not rows and b[0][2][0] is new_m
    rows = c.groups[0].buckets

    old_state = opt.state                                           # load_state_dict: a new state mapping
    opt.state = defaultdict(dict, {p: dict(st) for p, st in old_state.items()})
    c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is not rows and c.groups[0].state is opt.state
    rows = c.groups[0].buckets

    group["params"] = list(group["params"])                         # amp swaps the list object of a group
    c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is not rows
    rows = c.groups[0].buckets

    extra = torch.nn.Parameter(torch.zeros(7))                      # a parameter appended to the same list
    extra.grad = torch.ones(7)
    group["params"].append(extra)
    b = c.buckets(opt, 0, group, init_state)
    assert c.groups[0].buckets is not rows and len(b[0][0]) == 6 and "exp_avg" in opt.state[extra]


def test_buckets_by_dtype_and_reference_error_messages(GroupCache):
    c = GroupCache("FusedLAMB", ("exp_avg", "exp_avg_sq"))
    opt, group, params = make(6, (torch.float32, torch.float16, torch.bfloat16))
    b = c.buckets(opt, 0, group, init_state)
    assert [blk[1][0].dtype for blk in b] == [torch.float32, torch.float16, torch.bfloat16]
    assert [len(blk[0]) for blk in b] == [2, 2, 2]
    bad = torch.nn.Parameter(torch.zeros(4, dtype=torch.float64))
    bad.grad = torch.zeros(4, dtype=torch.float64)
    group["params"].append(bad)
    with pytest.raises(RuntimeError, match="FusedLAMB only support fp16, bfloat16 and fp32"):
        c.buckets(opt, 0, group, init_state)
    group["params"].pop()
    sp = torch.nn.Parameter(torch.zeros(4, 4))
    sp.grad = torch.zeros(4, 4).to_sparse()
    group["params"].append(sp)
    with pytest.raises(RuntimeError, match="FusedLAMB does not support sparse gradients"):
        c.buckets(opt, 0, group, init_state)


def test_no_gradients_at_all_gives_no_buckets(GroupCache):
    c = GroupCache("FusedAdam", ("exp_avg", "exp_avg_sq"))
    opt, group, params = make()
    for p in params:
        p.grad = None
    assert c.buckets(opt, 0, group, init_state) == []
    assert c.buckets(opt, 0, group, init_state) == []               # and the (empty) memo is consistent
    params[1].grad = torch.ones_like(params[1])
    b = c.buckets(opt, 0, group, init_state)
    assert len(b) == 1 and len(b[0][1]) == 1 and b[0][1][0] is params[1]


def test_native_row_check_codes():
    """csrc/packer.cpp: check_rows -- the one native loop that decides whether a prepared [g, p, s0, s1] handle may be
    reused: 0 = gradients at the cached addresses and state tensors unchanged, 1 = a gradient moved / vanished / changed
    type, 2 = a state tensor was replaced.  Works on any device's tensors: exercised here on CPU."""
    import numpy as np
    import pytest
    import torch
    import _b200mt as mt
    if mt.check_rows is None:
        pytest.skip("native packer not built")
    ps = [torch.nn.Parameter(torch.randn(n)) for n in (5, 64, 1000)]
    for p in ps:
        p.grad = torch.randn_like(p)
    states = [{"exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p), "other": 1} for p in ps]
    rows = [(p, st, st["exp_avg"], st["exp_avg_sq"]) for p, st in zip(ps, states)]
    ptrs = np.array([p.grad.data_ptr() for p in ps], dtype=np.uint64)
    f32 = mt.F32
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32) == 0
    ps[1].grad.add_(1.0)                                         # values change, addresses do not
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32) == 0
    old = ps[1].grad
    ps[1].grad = torch.randn_like(ps[1])                         # a gradient somewhere else
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32) == 1
    ps[1].grad = None                                            # ... or gone
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32) == 1
    ps[1].grad = old
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32) == 0
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, mt.F16) == 1     # cached dtype no longer matches
    states[2]["exp_avg_sq"] = torch.zeros_like(ps[2])            # a state tensor replaced by the user
    assert mt.check_rows(rows, "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32) == 2
    with pytest.raises(RuntimeError, match="malformed"):
        mt.check_rows([(ps[0], states[0], None)], "exp_avg", "exp_avg_sq", ptrs.ctypes.data, f32)
