"""CPU, world_size 2 over gloo: the host-side logic of DistributedFusedAdam -- flat layout, shard
fragments, reduce-scatter / all-gather plumbing, checkpoint shards.  There is no CPU kernel in the
product, so THIS TEST substitutes the three kernel modules by doubles that call the oracle
(tests may; the product never does), and checks the sharded result against unsharded AdamW on the
rank-averaged gradients, as tests/L0/run_optimizers/test_dist_adam.py does on GPUs."""
import os
import sys
import types

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_layout_and_fragments_cover_every_element_once():
    sys.path.insert(0, os.path.join(ROOT, "rocm-apex_b200"))
    from apex.contrib.optimizers.distributed_fused_adam import build_layout, shard_fragments
    numels = [1000, 64, 7, 300000, 1, 65536]
    for W, K in [(1, 1), (2, 1), (2, 4), (8, 16), (4, 3)]:
        L = build_layout(numels, W, K)
        assert L["total"] % (256 * K * W) == 0 and L["chunk"] * K == L["total"] and L["shard"] * W == L["chunk"]
        assert all(o % 64 == 0 for o in L["offsets"])
        seen = [torch.zeros(n, dtype=torch.int32) for n in numels]
        for r in range(W):
            for k, frags in enumerate(shard_fragments(L, numels, r)):
                used = 0
                for (pi, poff, soff, n) in frags:
                    seen[pi][poff:poff + n] += 1
                    assert 0 <= soff and soff + n <= L["shard"]
                    # same element, same flat position
                    assert L["offsets"][pi] + poff == k * L["chunk"] + r * L["shard"] + soff
                    used += n
        assert all(bool((s == 1).all()) for s in seen)


def _install_kernel_doubles():
    from oracle import mt_oracle as O

    def flag_list(t):
        class F(list):
            pass
        f = F([int(t.item())])
        return f

    amp_c = types.ModuleType("amp_C")

    def multi_tensor_scale(chunk, noop, lists, scale):
        f = flag_list(noop); O.multi_tensor_scale(f, lists, scale); noop.fill_(f[0])

    def multi_tensor_l2norm(chunk, noop, lists, per_tensor=False):
        f = flag_list(noop)
        n, p = O.multi_tensor_l2norm(chunk, f, lists, bool(per_tensor))
        noop.fill_(f[0])
        return torch.tensor([float(n)]), (torch.from_numpy(p) if p is not None else torch.empty(0))

    amp_c.multi_tensor_scale, amp_c.multi_tensor_l2norm = multi_tensor_scale, multi_tensor_l2norm
    dac = types.ModuleType("distributed_adam_cuda")

    def multi_tensor_fused_adam(chunk, noop, lists, b1, b2, bc, eps, wd, lr, grad_scale, step, mode):
        O.multi_tensor_fused_adam(lists, b1.tolist(), b2.tolist(), bc.tolist(), eps.tolist(), wd.tolist(), lr, grad_scale, step, mode)

    dac.multi_tensor_fused_adam = multi_tensor_fused_adam
    fac = types.ModuleType("fused_adam_cuda")

    def maybe_cast_mt(chunk, flag, lists):
        O.maybe_cast_mt([int(flag.item())], lists)

    fac.maybe_cast_mt = maybe_cast_mt
    sys.modules["amp_C"], sys.modules["distributed_adam_cuda"], sys.modules["fused_adam_cuda"] = amp_c, dac, fac


def _worker(rank, world, port, K, dtype_name, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401  (real launch shim first, so multi_tensor_applier.available is decided by the real library)
        _install_kernel_doubles()
        from apex.contrib.optimizers.distributed_fused_adam import DistributedFusedAdam
        dtype = getattr(torch, dtype_name)
        shapes = [(300, 7), (64,), (5,), (1000, 3), (1,)]
        g0 = torch.Generator().manual_seed(7)
        init = [torch.randn(s, generator=g0) * 0.1 for s in shapes]
        params = [torch.nn.Parameter(t.to(dtype)) for t in init]
        opt = DistributedFusedAdam([{"params": params[:3], "weight_decay": 0.01}, {"params": params[3:], "weight_decay": 0.0}],
                                   lr=1e-2, betas=(0.9, 0.99), eps=1e-8, dwu_num_blocks=K, dwu_num_chunks=1,
                                   compute_L2_grad_norm=True, max_grad_norm=0.0, overlap_reductions=False)
        opt.set_global_scale(8.0)
        # unsharded reference: AdamW on fp32 masters with the rank-averaged gradient
        masters = [p.detach().float().clone().requires_grad_() for p in params]
        ref = torch.optim.AdamW([{"params": masters[:3], "weight_decay": 0.01}, {"params": masters[3:], "weight_decay": 0.0}],
                                lr=1e-2, betas=(0.9, 0.99), eps=1e-8)
        for it in range(4):
            per_rank = []
            for r in range(world):
                gr = torch.Generator().manual_seed(100 * it + r)
                per_rank.append([(torch.randn(s, generator=gr) * 8.0).to(dtype) for s in shapes])
            for p, g in zip(params, per_rank[rank]):
                p.grad = g.clone()
            opt.complete_reductions()
            # averaged, still-scaled gradient as the reduce-scatter produced it (each term rounded to dtype)
            avg = [sum((gs[i].float() / world).to(dtype).float() for gs in per_rank).to(dtype).float() for i in range(len(shapes))]
            want_norm = torch.sqrt(sum((a * a).sum() for a in avg))
            assert abs(opt.L2_grad_norm - float(want_norm)) <= 1e-3 * float(want_norm)
            opt.step()
            for m, a in zip(masters, avg):
                m.grad = a / 8.0
            ref.step()
            tol = 1e-5 if dtype == torch.float32 else 1e-2
            for p, m in zip(params, masters):
                assert torch.allclose(p.detach().float(), m.detach().to(dtype).float(), rtol=tol, atol=tol), (it, rank)
        sd = opt.state_dict()
        assert sd["step"] == 4 and sd["fp32_p"].numel() == opt._layout["shard"] * K
        # resume: a fresh optimizer loaded from this rank's shard continues identically
        params2 = [torch.nn.Parameter(p.detach().clone()) for p in params]
        opt2 = DistributedFusedAdam([{"params": params2[:3], "weight_decay": 0.01}, {"params": params2[3:], "weight_decay": 0.0}],
                                    lr=1e-2, betas=(0.9, 0.99), eps=1e-8, dwu_num_blocks=K, dwu_num_chunks=1, overlap_reductions=False)
        opt2.load_state_dict({k: (v.clone() if torch.is_tensor(v) else v) for k, v in sd.items()})
        opt2.set_global_scale(8.0)
        gr = torch.Generator().manual_seed(999 + rank)
        gnew = [(torch.randn(s, generator=gr) * 8.0).to(dtype) for s in shapes]
        for p, p2, g in zip(params, params2, gnew):
            p.grad, p2.grad = g.clone(), g.clone()
        opt.step(); opt2.step()
        for p, p2 in zip(params, params2):
            assert torch.equal(p.detach(), p2.detach())
        out.put((rank, "ok"))
    except Exception as e:  # surface the failure to the parent
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _worker_e5m2(rank, world, port, K, dtype_name, out):
    """e5m2_allgather=True beside a plain optimizer on the same gradients: identical fp32 shards, model
    parameters = the e5m2 rounding of the masters (reference distributed_fused_adam.py:198,206,584-591)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        _install_kernel_doubles()
        from apex.contrib.optimizers.distributed_fused_adam import DistributedFusedAdam
        from oracle import mt_oracle as O
        dtype = getattr(torch, dtype_name)
        shapes = [(300, 7), (64,), (5,), (1000, 3), (1,)]
        g0 = torch.Generator().manual_seed(7)
        init = [torch.randn(s, generator=g0) * 0.1 for s in shapes]
        pa = [torch.nn.Parameter(t.to(dtype)) for t in init]
        pb = [torch.nn.Parameter(t.to(dtype)) for t in init]
        kw = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_num_blocks=K, dwu_num_chunks=1,
                  overlap_reductions=False)
        oa = DistributedFusedAdam(pa, e5m2_allgather=True, **kw)
        ob = DistributedFusedAdam(pb, **kw)
        with pytest.raises(ValueError):
            DistributedFusedAdam(pb, e5m2_allgather=True, zero_copy_params=True, **kw)
        for it in range(3):
            gr = torch.Generator().manual_seed(100 * it + rank)
            grads = [(torch.randn(s, generator=gr) * 4.0).to(dtype) for s in shapes]
            for a, b, g in zip(pa, pb, grads):
                a.grad, b.grad = g.clone(), g.clone()
            for o in (oa, ob):
                o.set_global_scale(4.0)
                o.step()
            assert oa._new_params.dtype == torch.uint8 and oa._fp16_p.dtype == torch.uint8
            assert torch.equal(oa._fp32_p, ob._fp32_p) and torch.equal(oa._fp32_m, ob._fp32_m)
            # the gathered bytes are the packing of the masters: unpack(pack(.)) of the plain run's full-precision
            # masters is not available on this rank for other shards, so compare through this rank's own shard ...
            shard = oa._layout["shard"]
            for k in range(K):
                lo = (k * world + rank) * shard
                assert torch.equal(oa._new_params[lo:lo + shard],
                                   torch.from_numpy(O.pack_e5m2(oa._fp32_p[k * shard:(k + 1) * shard].numpy())))
            # ... and every model parameter is an e5m2 value within e5m2 rounding (1/8 relative) of the plain one
            for a, b in zip(pa, pb):
                h = a.detach().to(torch.float16).view(torch.int16)
                assert bool(((h & 0xFF) == 0).all())
                d = (a.detach().float() - b.detach().float()).abs()
                assert bool((d <= 0.13 * b.detach().float().abs() + 2e-5).all())
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _worker_overlap(rank, world, port, K, dtype_name, out):
    """overlap_reductions=True: gradient hooks flush pipeline chunks during backward, last chunk first
    (reference :158-164,381-395,489-502); the result is bit-identical to reducing everything afterwards."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        _install_kernel_doubles()
        from apex.contrib.optimizers.distributed_fused_adam import DistributedFusedAdam
        dtype = getattr(torch, dtype_name)
        shapes = [(300, 7), (64,), (5,), (1000, 3), (1,), (700,)]
        g0 = torch.Generator().manual_seed(7 + rank)             # ranks start differently: rank 0's values must win
        init = [torch.randn(s, generator=g0) * 0.1 for s in shapes]
        pa = [torch.nn.Parameter(t.to(dtype)) for t in init]
        pb = [torch.nn.Parameter(t.to(dtype)) for t in init]
        kw = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_num_blocks=K, dwu_num_chunks=1)
        oa = DistributedFusedAdam(pa, overlap_reductions=True, **kw)
        ob = DistributedFusedAdam(pb, overlap_reductions=False, **kw)
        oa._init_everything(); ob._init_everything()             # hooks exist from the first backward on
        first = [torch.zeros_like(p.data) for p in pa]
        for f, p in zip(first, pa):
            f.copy_(p.data)
            dist.broadcast(f, src=0)
            assert torch.equal(f, p.data)                        # every rank holds rank 0's parameters
        assert len(oa._hooks) == len(pa) and len(ob._hooks) == 0
        launched = []
        orig = oa._reduce_scatter_chunk
        oa._reduce_scatter_chunk = lambda k: (launched.append(k), orig(k))[1]
        for it in range(4):
            gr = torch.Generator().manual_seed(100 * it + rank)
            xs = [torch.randn(s, generator=gr).to(dtype) for s in shapes]
            skip = 1 if it == 2 else None                        # iteration 2: parameter 1 takes no part in the loss
            last = it == 3                                       # iteration 3: set_last_step(True) -> no overlap
            oa.set_last_step(last)
            for params, opt in ((pa, oa), (pb, ob)):
                for p in params:
                    p.grad = None
                loss = sum((p.float() * x.float()).sum() for i, (p, x) in enumerate(zip(params, xs)) if i != skip)
                del launched[:]
                loss.backward()
                if opt is oa:
                    if last:
                        assert launched == [] and oa._next_chunk == K - 1
                    elif skip is None:
                        assert launched == list(range(K - 1, -1, -1)) and oa._next_chunk == -1
                    else:
                        # the chunk(s) holding parameter 1 and every chunk below stay pending
                        k1 = min(k for k in range(K) if 1 in oa._chunk_params[k])
                        assert oa._next_chunk >= k1 and launched == list(range(K - 1, oa._next_chunk, -1))
                opt.complete_reductions()
                if opt is oa:
                    assert sorted(launched) == list(range(K)) and oa._next_chunk == K - 1
                opt.step()
            for a, b in zip(pa, pb):
                assert torch.equal(a.detach(), b.detach()), (it, rank)
            assert torch.equal(oa._fp32_p, ob._fp32_p) and torch.equal(oa._fp32_v, ob._fp32_v)
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _worker_groups(rank, world, port, K, gs, out):
    """dwu_group_size < world (reference :115-117,328-345,411-418): shards inside groups of `gs` ranks, state
    replicated across groups, gradients reduce-scattered in the group and all-reduced across groups."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        _install_kernel_doubles()
        from apex.contrib.optimizers.distributed_fused_adam import DistributedFusedAdam
        shapes = [(300, 7), (64,), (5,), (1000, 3), (1,)]
        g0 = torch.Generator().manual_seed(7)
        init = [torch.randn(s, generator=g0) * 0.1 for s in shapes]
        params = [torch.nn.Parameter(t.clone()) for t in init]
        opt = DistributedFusedAdam(params, lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_group_size=gs,
                                   dwu_num_blocks=K, dwu_num_chunks=1, compute_L2_grad_norm=True, overlap_reductions=False)
        opt.set_global_scale(8.0)
        assert opt._world_size == gs and opt._rank == rank % gs and opt._grad_world == world
        masters = [t.clone().requires_grad_() for t in init]
        ref = torch.optim.AdamW(masters, lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01)
        for it in range(3):
            per_rank = []
            for r in range(world):
                gr = torch.Generator().manual_seed(100 * it + r)
                per_rank.append([torch.randn(s, generator=gr) * 8.0 for s in shapes])
            for p, g in zip(params, per_rank[rank]):
                p.grad = g.clone()
            opt.complete_reductions()
            avg = [sum(gs_[i] / world for gs_ in per_rank) for i in range(len(shapes))]
            want_norm = torch.sqrt(sum((a * a).sum() for a in avg))
            assert abs(opt.L2_grad_norm - float(want_norm)) <= 1e-4 * float(want_norm)
            opt.step()
            for m, a in zip(masters, avg):
                m.grad = a / 8.0
            ref.step()
            for p, m in zip(params, masters):
                assert torch.allclose(p.detach(), m.detach(), rtol=1e-5, atol=1e-6), (it, rank)
        # state is sharded inside a group and replicated across groups
        assert opt._fp32_p.numel() == opt._layout["shard"] * K and opt._layout["world_size"] == gs
        gathered = [torch.zeros_like(opt._fp32_m) for _ in range(world)]
        dist.all_gather(gathered, opt._fp32_m)
        for r in range(world):
            assert torch.equal(gathered[r], gathered[r % gs])
        assert not torch.equal(gathered[0], gathered[1])
        with pytest.raises(ValueError):
            DistributedFusedAdam(params, dwu_group_size=3)
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_rank_groups_world4_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_groups, args=(r, 4, 29680, 2, 2, q)) for r in range(4)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.parametrize("K,dtype_name", [(3, "float32"), (4, "bfloat16")])
def test_overlapped_reductions_world2_gloo(K, dtype_name):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29660 + K
    procs = [ctx.Process(target=_worker_overlap, args=(r, 2, port, K, dtype_name, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.parametrize("K,dtype_name", [(2, "float16"), (1, "bfloat16")])
def test_e5m2_allgather_world2_gloo(K, dtype_name):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29640 + K
    procs = [ctx.Process(target=_worker_e5m2, args=(r, 2, port, K, dtype_name, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.parametrize("K,dtype_name", [(1, "bfloat16"), (3, "float32")])
def test_sharded_adam_matches_unsharded_world2_gloo(K, dtype_name):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + K + (0 if dtype_name == "float32" else 10)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, K, dtype_name, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def _worker_parity_tool(rank, world, port, K, out):
    """tools/multi_gpu_parity.sharded_adam_parity (the check bench.py's N > 1 arm runs on the GPUs) on gloo with the
    oracle doubles: it must pass on a correct optimizer and fail when one shard element is corrupted."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT, os.path.join(ROOT, "tools")):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        _install_kernel_doubles()
        import multi_gpu_parity as mgp
        from apex.contrib.optimizers.distributed_fused_adam import DistributedFusedAdam
        shapes = [(300, 7), (64,), (5,), (1000, 3), (1,), (4096,)]

        def build():
            g0 = torch.Generator().manual_seed(7)
            params = [torch.nn.Parameter((torch.randn(s, generator=g0) * 0.1).bfloat16()) for s in shapes]
            opt = DistributedFusedAdam(params, lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_num_blocks=K,
                                       dwu_num_chunks=1, predivide=False, overlap_reductions=False, zero_copy_params=True)
            opt.set_global_scale(float(world))
            flat = opt.attach_flat_grads()
            flat.copy_(torch.randn(flat.numel(), generator=torch.Generator().manual_seed(100 + rank)))
            return opt

        res, ok = mgp.sharded_adam_parity(build(), sample_elems=512)
        assert ok, res
        assert res["master_p_m_v_max_rel"] == 0.0 and res["copy_vs_unsharded_max_ulp"] == 0
        assert res["elements_checked"] > 0 and res["reduced_grad_vs_rn_fp32_sum_max_ulp"] <= 1
        # a wrong update on one rank must be caught: corrupt the optimizer's step kernel for rank 1
        opt = build()
        if rank == 1:
            real = opt._dist_adam.multi_tensor_fused_adam

            def broken(chunk, noop, lists, *a):
                real(chunk, noop, lists, *a)
                lists[0][0][0] += 1.0          # first element of the shard's first fragment: inside the head sample

            opt._dist_adam = types.SimpleNamespace(multi_tensor_fused_adam=broken)
        res, ok = mgp.sharded_adam_parity(opt, sample_elems=512)
        assert not ok and res["master_p_m_v_max_rel"] > 1e-6, res
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("K", [1, 4])
def test_bench_parity_tool_world2_gloo(K):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_parity_tool, args=(r, 2, 29640 + K, K, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def _worker_overflow_then_clean(rank, world, port, out):
    """One iteration with a non-finite gradient, then a clean one: the overflow flag of the flatten pass gates the
    gathered-parameters -> model-parameters copy for THAT iteration only (the reference clears it before every flatten,
    apex/contrib/optimizers/distributed_fused_adam.py:481); on the clean iteration the model parameters must move."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        _install_kernel_doubles()
        from apex.contrib.optimizers.distributed_fused_adam import DistributedFusedAdam
        shapes = [(300, 7), (64,), (1000, 3)]
        g0 = torch.Generator().manual_seed(7)
        params = [torch.nn.Parameter(torch.randn(s, generator=g0) * 0.1) for s in shapes]
        opt = DistributedFusedAdam(params, lr=1e-2, betas=(0.9, 0.99), eps=1e-8, dwu_num_blocks=2, dwu_num_chunks=1,
                                   overlap_reductions=False)
        opt.set_global_scale(1.0)
        for it in range(3):
            gr = torch.Generator().manual_seed(100 * it + rank)
            grads = [torch.randn(s, generator=gr) for s in shapes]
            if it == 0 and rank == 1:
                grads[1][3] = float("inf")
            for p, g in zip(params, grads):
                p.grad = g
            before = [p.detach().clone() for p in params]
            opt.complete_reductions()
            opt.step()
            if it == 0:
                if rank == 1:       # the rank that saw the inf skips its copy for this iteration, like the reference
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


def test_overflow_flag_is_cleared_every_iteration_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_overflow_then_clean, args=(r, 2, 29660, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)
