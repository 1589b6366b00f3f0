"""GPU (needs >= 2 devices, NCCL): DistributedFusedAdam with the real kernels, sharded over 2 ranks,
against the unsharded mixed-precision Adam on the rank-averaged gradient -- the comparison
tests/L0/run_optimizers/test_dist_adam.py makes (DistributedFusedAdam vs FusedAdam, atol 1e-3)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, K, zero_copy, fused, out, e5m2=False, group_size=0):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import amp_C
        from apex.contrib.optimizers import DistributedFusedAdam
        shapes = [(1024, 300), (4096,), (5,), (3000, 77), (1,), (65536,)]
        g0 = torch.Generator().manual_seed(7)
        init = [torch.randn(s, generator=g0) * 0.1 for s in shapes]
        params = [torch.nn.Parameter(t.to(dev).bfloat16()) for t in init]
        opt = DistributedFusedAdam(params, lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_num_blocks=K,
                                   dwu_num_chunks=1, compute_L2_grad_norm=not fused, overlap_reductions=False,
                                   zero_copy_params=zero_copy, fused_collectives=fused, e5m2_allgather=e5m2,
                                   dwu_group_size=group_size)
        opt.set_global_scale(8.0)
        import fused_adam_cuda
        # unsharded check: our own mixed-precision multi_tensor_adam [g, p32, m32, v32, p_copy] on the averaged grads
        m32 = [t.to(dev).bfloat16().float() for t in init]
        mm = [torch.zeros_like(t) for t in m32]
        vv = [torch.zeros_like(t) for t in m32]
        pc = [torch.zeros_like(t, dtype=torch.bfloat16) for t in m32]
        flag = torch.zeros(1, dtype=torch.int32, device=dev)
        for it in range(1, 5):
            per_rank = []
            for r in range(world):
                gr = torch.Generator().manual_seed(100 * it + r)
                per_rank.append([(torch.randn(s, generator=gr) * 8.0).to(dev).bfloat16() for s in shapes])
            for p, g in zip(params, per_rank[rank]):
                p.grad = g.clone()
            opt.complete_reductions()
            opt.step()
            avg = [sum((gs[i].float() / world).bfloat16().float() for gs in per_rank).bfloat16() for i in range(len(shapes))]
            amp_C.multi_tensor_adam(65536, flag, [[a.float() / 8.0 for a in avg], m32, mm, vv, pc], 1e-2, 0.9, 0.99, 1e-8, it, 1, 1, 0.01)
            want_norm = torch.sqrt(sum((a.float() ** 2).sum() for a in avg))
            if not fused:
                assert abs(opt.L2_grad_norm - float(want_norm)) <= 1e-3 * float(want_norm)
            if e5m2:
                # model parameters = e5m2 rounding of the fp32 masters (bytes travelled, maybe_cast_mt unpacked them)
                assert opt._new_params.dtype == torch.uint8
                for p, m in zip(params, m32):
                    byts = torch.zeros(m.numel(), dtype=torch.uint8, device=dev)
                    want = torch.zeros(m.numel(), dtype=torch.bfloat16, device=dev)
                    fused_adam_cuda.maybe_cast_mt(65536, flag, [[m.reshape(-1)], [byts]])
                    fused_adam_cuda.maybe_cast_mt(65536, flag, [[byts], [want]])
                    a, b = p.detach().float().reshape(-1), want.float()
                    assert bool(((p.detach().reshape(-1).to(torch.float16).view(torch.int16) & 0xFF) == 0).all())
                    assert float((a == b).float().mean()) >= (0.98 if fused else 0.999), (it, rank)
                    assert bool(((a - m.reshape(-1)).abs() <= 0.126 * m.reshape(-1).abs() + 3e-4).all())
                continue
            for pi, (p, c) in enumerate(zip(params, pc)):
                # bf16 parameters: bit-exact expected, one bf16 ulp allowed where NCCL's summation order differs
                a, b = p.detach().float().reshape(-1), c.float().reshape(-1)
                # fused_collectives: the gradient sum is formed inside the NVSwitch (multimem.ld_reduce) and
                # comes back rounded to bf16 by the switch, which is not always the round-to-nearest-even of the
                # fp32 sum this check uses: a gradient may differ by one bf16 ulp (2^-8 relative), and the Adam
                # update (|u| <~ 3, lr 1e-2) turns that into at most ~lr*3*2^-7 = 2.4e-4 absolute on p
                atol = 2.5e-4 if fused else 1e-6
                bad = (a - b).abs() > (8e-3 * b.abs() + atol)
                if bad.any():
                    idx = torch.nonzero(bad).reshape(-1)
                    raise AssertionError("it %d rank %d param %d: %d/%d differ, first idx %d..last %d, got %r want %r" % (
                        it, rank, pi, int(bad.sum()), a.numel(), int(idx[0]), int(idx[-1]), float(a[idx[0]]), float(b[idx[0]])))
        # master shards reassemble to the unsharded fp32 masters
        L = opt._layout
        full = torch.zeros(L["total"], device=dev)
        gathered = [torch.zeros_like(opt._fp32_p) for _ in range(world)]
        dist.all_gather(gathered, opt._fp32_p)
        for r in range(world):
            for k in range(K):
                lo = k * L["chunk"] + (r % L["world_size"]) * L["shard"]     # rank groups replicate the shards
                full[lo:lo + L["shard"]] = gathered[r][k * L["shard"]:(k + 1) * L["shard"]]
        for off, t in zip(L["offsets"], m32):
            got = full[off:off + t.numel()].view_as(t)
            assert torch.allclose(got, t, rtol=1e-5, atol=2.5e-4 if fused else 1e-6)
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("K,zero_copy,fused", [(1, False, False), (4, True, False), (1, False, True), (1, True, True),
                                               (1, False, "multimem"), (1, True, "p2p")])
def test_sharded_adam_matches_unsharded_world2_nccl(K, zero_copy, fused):
    """fused: False = NCCL pipeline; True = one kernel, form picked by world size (two ranks: peer-memory loads and stores);
    "multimem" / "p2p" force the in-switch or the peer-memory form."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + K + (20 if fused else 0) + (40 if zero_copy else 0) + (5 if isinstance(fused, str) else 0)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, K, zero_copy, fused, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_rank_groups_of_one_world2_nccl():
    """dwu_group_size=1 on 2 ranks: every rank is its own shard group, the gradient 'shard' (everything) is all-reduced
    across the two groups on the communication stream -- the cross-group leg of the multi-node layout over NCCL."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29790, 2, False, False, q, False, 1)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def _worker_overlap(rank, world, port, K, out):
    """Reductions started by gradient hooks during a real backward pass (side stream, NCCL) give the same bits as
    reducing everything in complete_reductions()."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from apex.contrib.optimizers import DistributedFusedAdam
        torch.manual_seed(11)
        def net():
            return torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.GELU(), torch.nn.Linear(512, 512), torch.nn.GELU(),
                                       torch.nn.Linear(512, 64)).to(dev).bfloat16()
        ma = net()
        mb = net()
        mb.load_state_dict(ma.state_dict())
        kw = dict(lr=1e-3, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_num_blocks=K, dwu_num_chunks=1)
        oa = DistributedFusedAdam(ma.parameters(), overlap_reductions=True, **kw)
        ob = DistributedFusedAdam(mb.parameters(), overlap_reductions=False, **kw)
        oa._init_everything(); ob._init_everything()
        flushed_in_backward = 0
        for it in range(4):
            x = torch.randn(64, 256, generator=torch.Generator().manual_seed(100 * it + rank)).to(dev).bfloat16()
            for model, opt in ((ma, oa), (mb, ob)):
                for p in model.parameters():
                    p.grad = None
                model(x).float().pow(2).mean().backward()
                if opt is oa:
                    flushed_in_backward += (K - 1) - oa._next_chunk
                opt.complete_reductions()
                opt.step()
            for a, b in zip(ma.parameters(), mb.parameters()):
                assert torch.equal(a.detach(), b.detach()), (it, rank)
        assert flushed_in_backward == 4 * K          # every chunk of every iteration left during backward
        assert torch.equal(oa._fp32_p, ob._fp32_p) and torch.equal(oa._fp32_m, ob._fp32_m)
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_overlapped_reductions_world2_nccl():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_overlap, args=(r, 2, 29795, 4, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("K,fused", [(2, False), (1, True)])
def test_e5m2_allgather_world2_nccl(K, fused):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29780 + K + (20 if fused else 0)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, K, False, fused, q, True)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def _spawn(target, world, args):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=target, args=(r, world) + tuple(args) + (q,)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.parametrize("K,zero_copy", [(1, False), (4, True)])
def test_sharded_adam_world1_nccl_one_gpu(K, zero_copy):
    """The same comparison on ONE GPU (a 1-rank NCCL job): no cross-GPU traffic, but the class, the flat layout, the
    fragment lists, reduce_scatter_tensor / all_gather_into_tensor and every kernel of the step run for real on a box
    that has a single device."""
    _spawn(_worker_w1, 1, (29600 + K, K, zero_copy))


def _worker_w1(rank, world, port, K, zero_copy, out):
    _worker(rank, world, port, K, zero_copy, False, out)


def _worker_overflow(rank, world, port, out):
    """An iteration with an inf gradient, then clean ones, with the real kernels: the flatten pass's overflow flag gates
    the gathered -> model copy for that iteration only (reference distributed_fused_adam.py:481 clears it per flatten)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from apex.contrib.optimizers import DistributedFusedAdam
        shapes = [(300, 7), (64,), (1000, 3)]
        g0 = torch.Generator().manual_seed(7)
        params = [torch.nn.Parameter((torch.randn(s, generator=g0) * 0.1).to(dev).half()) for s in shapes]
        opt = DistributedFusedAdam(params, lr=1e-2, betas=(0.9, 0.99), eps=1e-8, dwu_num_blocks=2, dwu_num_chunks=1,
                                   overlap_reductions=False)
        opt.set_global_scale(1.0)
        for it in range(3):
            gr = torch.Generator().manual_seed(100 * it + rank)
            grads = [torch.randn(s, generator=gr).to(dev).half() for s in shapes]
            if it == 0 and rank == world - 1:
                grads[1][3] = float("inf")
            for p, g in zip(params, grads):
                p.grad = g
            before = [p.detach().clone() for p in params]
            opt.complete_reductions()
            opt.step()
            torch.cuda.synchronize()
            if it == 0:
                if rank == world - 1:
                    assert int(opt._overflow_buf) == 1 and all(torch.equal(a, b.detach()) for a, b in zip(before, params))
            else:
                assert int(opt._overflow_buf) == 0, (it, rank)
                assert not torch.equal(before[0], params[0].detach()), "model parameters stopped following the masters"
                assert not torch.equal(before[2], params[2].detach())
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_overflow_flag_cleared_every_iteration_world1_nccl_one_gpu():
    _spawn(_worker_overflow, 1, (29610,))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_overflow_flag_cleared_every_iteration_world2_nccl():
    _spawn(_worker_overflow, 2, (29611,))
