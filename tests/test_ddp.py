"""apex.parallel.DistributedDataParallel / Reducer.
CPU (gloo, world 2): bucket logic, hooks and collectives, with amp_C.multi_tensor_scale replaced by an oracle
double inside the test.  GPU (NCCL, world 2): the real kernels on side streams.  Both check that after
backward every rank holds the mean of the per-rank gradients, for overlapped and delayed all-reduce, with
predivide, fp32 communication of fp16 gradients, retained buffers, and disable/enable (gradient accumulation)."""
import os
import sys
import types

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _paths():
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)


def _install_cpu_double():
    from oracle import mt_oracle as O
    amp_c = types.ModuleType("amp_C")

    def multi_tensor_scale(chunk, noop, lists, scale):
        f = [int(noop.item())]
        O.multi_tensor_scale(f, lists, scale)
        noop.fill_(f[0])

    amp_c.multi_tensor_scale = multi_tensor_scale
    sys.modules["amp_C"] = amp_c


class Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(32, 64)
        self.b = torch.nn.Linear(64, 64)
        self.c = torch.nn.Linear(64, 8)
        self.unused = torch.nn.Parameter(torch.zeros(5))      # never gets a gradient

    def forward(self, x):
        return self.c(torch.relu(self.b(torch.relu(self.a(x)))))


def _local_grads(model, x):
    for p in model.parameters():
        p.grad = None
    model(x).pow(2).mean().backward()
    return [None if p.grad is None else p.grad.detach().clone() for p in model.parameters()]


def _worker(rank, world, port, backend, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    _paths()
    cuda = backend == "nccl"
    dev = torch.device("cuda", rank) if cuda else torch.device("cpu")
    if cuda:
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        if not cuda:
            _install_cpu_double()
        from apex.parallel import DistributedDataParallel as DDP, Reducer
        import copy
        torch.manual_seed(100 + rank)                      # different init per rank: the constructor must broadcast rank 0's
        base = Net().to(dev)
        for kw in (dict(message_size=3000), dict(delay_allreduce=True), dict(message_size=1, gradient_predivide_factor=2.0),
                   dict(message_size=3000, retain_allreduce_buffers=True), dict(message_size=3000, gradient_average=False),
                   dict(message_size=2000, num_allreduce_streams=2)):
            model = copy.deepcopy(base)
            ddp = DDP(model, **kw)
            w0 = [p.detach().clone() for p in model.parameters()]
            gathered = [torch.zeros_like(w0[0]) for _ in range(world)]
            dist.all_gather(gathered, w0[0])
            assert all(torch.equal(g, gathered[0]) for g in gathered), "parameters were not broadcast"
            ref_model = copy.deepcopy(model)
            for it in range(3):                             # iteration 1 builds the buckets, 2 and 3 overlap
                xs = [torch.randn(16, 32, generator=torch.Generator().manual_seed(1000 * it + r)).to(dev) for r in range(world)]
                want = [_local_grads(ref_model, x) for x in xs]
                for p in model.parameters():
                    p.grad = None
                ddp(xs[rank]).pow(2).mean().backward()
                if cuda:
                    torch.cuda.synchronize()
                div = world if kw.get("gradient_average", True) else 1
                if kw.get("retain_allreduce_buffers"):
                    flat_all = torch.cat([b.flatten() for b in ddp.allreduce_buffers])
                    assert flat_all.numel() == sum(p.numel() for p in model.parameters())
                    assert len(ddp._buckets) >= 2
                    continue
                for i, p in enumerate(model.parameters()):
                    if want[0][i] is None:
                        assert p.grad is None or float(p.grad.abs().max()) == 0.0
                        continue
                    mean = sum(w[i] for w in want) / div
                    assert torch.allclose(p.grad, mean, rtol=1e-5, atol=1e-7), (kw, it, i)
            if "message_size" in kw and kw["message_size"] == 3000:
                assert len(ddp._buckets) >= 2
        # gradient accumulation: no communication while disabled, one reduction of the accumulated gradients after
        model = copy.deepcopy(base)
        ddp = DDP(model, message_size=3000)
        ref_model = copy.deepcopy(model)
        xs = [[torch.randn(16, 32, generator=torch.Generator().manual_seed(7 * k + r)).to(dev) for r in range(world)] for k in range(2)]
        ddp.disable_allreduce()
        ddp(xs[0][rank]).pow(2).mean().backward()
        local = [None if p.grad is None else p.grad.clone() for p in model.parameters()]
        want_local = _local_grads(ref_model, xs[0][rank])
        for a, b in zip(local, want_local):
            assert (a is None and b is None) or torch.allclose(a, b, rtol=1e-6, atol=1e-8)
        ddp.enable_allreduce()
        ddp(xs[1][rank]).pow(2).mean().backward()
        if cuda:
            torch.cuda.synchronize()
        acc = []
        for r in range(world):
            g0, g1 = _local_grads(ref_model, xs[0][r]), _local_grads(ref_model, xs[1][r])
            acc.append([None if a is None else a + b for a, b in zip(g0, g1)])
        for i, p in enumerate(model.parameters()):
            if acc[0][i] is not None:
                assert torch.allclose(p.grad, sum(a[i] for a in acc) / world, rtol=1e-5, atol=1e-7)
        # fp16 gradients communicated in fp32
        if cuda:
            half = copy.deepcopy(base).half()
            ddp = DDP(half, message_size=3000, allreduce_always_fp32=True)
            for it in range(2):
                for p in half.parameters():
                    p.grad = None
                ddp(torch.randn(16, 32, device=dev, dtype=torch.float16)).float().pow(2).mean().backward()
            torch.cuda.synchronize()
            g = half.a.weight.grad.float()
            gs = [torch.zeros_like(g) for _ in range(world)]
            dist.all_gather(gs, g)
            assert all(torch.equal(gs[0], x) for x in gs) and ddp._buckets[0].flat.dtype == torch.float32
        # Reducer: manual all-reduce of a module's gradients
        model = copy.deepcopy(base)
        red = Reducer(model)
        want = [_local_grads(model, torch.randn(16, 32, generator=torch.Generator().manual_seed(50 + r)).to(dev)) for r in range(world)]
        _local_grads(model, torch.randn(16, 32, generator=torch.Generator().manual_seed(50 + rank)).to(dev))
        red.reduce()
        for i, p in enumerate(model.parameters()):
            if want[0][i] is not None:
                assert torch.allclose(p.grad, sum(w[i] for w in want) / world, rtol=1e-5, atol=1e-7)
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _run(backend, port, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, backend, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def test_ddp_world2_gloo():
    _run("gloo", 29770)


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_ddp_world2_nccl():
    _run("nccl", 29771)


@pytest.mark.gpu
def test_ddp_world1_nccl_one_gpu():
    """1-rank NCCL job on one GPU: bucket building, the flat-bucket multi_tensor_scale gather / scatter launches on the
    side streams and the all_reduce call run for real; the expected gradient is the local one."""
    _run("nccl", 29772, world=1)


class TwoBranch(torch.nn.Module):
    """Two independent same-shaped layers: the order in which they are evaluated decides the order in which autograd
    produces their gradients (later nodes first)."""

    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(32, 32)
        self.b = torch.nn.Linear(32, 32)

    def forward(self, x, a_last=False, skip_b=False):
        if skip_b:
            return self.a(x).pow(2).mean()
        if a_last:
            yb = self.b(x); ya = self.a(x)
        else:
            ya = self.a(x); yb = self.b(x)
        return ya.pow(2).mean() + 2.0 * yb.pow(2).mean()


def _worker_order(rank, world, port, backend, out):
    """Ranks whose gradients arrive in different orders must still issue their all_reduce calls in bucket order (the old
    code launched buckets as they filled: same-sized buckets were silently mixed); trigger parameters cut the buckets;
    a bucket that never fills in an overlapped iteration is an error, as in the reference's epilogue."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    _paths()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        _install_cpu_double()
        from apex.parallel import DistributedDataParallel as DDP
        import copy
        torch.manual_seed(5)
        base = TwoBranch()
        model = copy.deepcopy(base)
        ddp = DDP(model, message_size=1)                    # one bucket per parameter
        for it in range(4):
            xs = [torch.randn(8, 32, generator=torch.Generator().manual_seed(10 * it + r)) for r in range(world)]
            want = []
            for r in range(world):
                ref = copy.deepcopy(base)
                ref(xs[r]).backward()
                want.append([p.grad.clone() for p in ref.parameters()])
            for p in model.parameters():
                p.grad = None
            ddp(xs[rank], a_last=(rank == 1)).backward()    # rank 1 produces a's gradients first, rank 0 b's
            for i, p in enumerate(model.parameters()):
                mean = sum(w[i] for w in want) / world
                assert torch.allclose(p.grad, mean, rtol=1e-5, atol=1e-7), (it, rank, i)
        assert len(ddp._buckets) == 4 and ddp._num_regular == 4
        # a regular bucket that does not fill: refused in the epilogue (every rank skips b here, so nobody hangs)
        for p in model.parameters():
            p.grad = None
        try:
            ddp(torch.randn(8, 32), skip_b=True).backward()
            raise AssertionError("an unfilled overlapped bucket went unnoticed")
        except RuntimeError as e:
            assert "next_bucket" in str(e)
        # trigger parameters: buckets end where a trigger's gradient arrives (reference :373-376)
        model = copy.deepcopy(base)
        ddp = DDP(model, allreduce_trigger_params=[model.b.weight])
        for it in range(3):
            xs = [torch.randn(8, 32, generator=torch.Generator().manual_seed(70 * it + r)) for r in range(world)]
            want = []
            for r in range(world):
                ref = copy.deepcopy(base)
                ref(xs[r]).backward()
                want.append([p.grad.clone() for p in ref.parameters()])
            for p in model.parameters():
                p.grad = None
            ddp(xs[rank]).backward()
            for i, p in enumerate(model.parameters()):
                assert torch.allclose(p.grad, sum(w[i] for w in want) / world, rtol=1e-5, atol=1e-7), (it, rank, i)
        # rank 0's iteration-1 order is b.bias/b.weight (b evaluated last), then a.*: cut after b.weight -> 2 buckets
        assert len(ddp._buckets) == 2 and any(p is model.b.weight for p in ddp._buckets[0].params)
        assert all(p is not model.b.weight for p in ddp._buckets[1].params)
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_ddp_bucket_order_and_triggers_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_order, args=(r, 2, 29773, "gloo", q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)
