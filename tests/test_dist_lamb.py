"""DistributedFusedLAMB.
CPU (gloo, world 2): host logic with the kernel modules replaced by oracle doubles inside the test.
GPU (NCCL, world 2, needs 2 devices): real kernels, against the unsharded FusedLAMB on the averaged gradients.
GPU (1 device): the two sharded-LAMB kernels against their CPU restatement (oracle/dist_lamb_oracle.py)."""
import os
import sys
import types

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHAPES = [(300, 7), (64,), (5,), (1000, 3), (1,), (4096,)]


def _paths():
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)


def _install_cpu_doubles():
    from oracle import mt_oracle as O
    from oracle import dist_lamb_oracle as DL
    amp_c = types.ModuleType("amp_C")

    def multi_tensor_scale(chunk, noop, lists, scale):
        f = [int(noop.item())]; O.multi_tensor_scale(f, lists, scale); noop.fill_(f[0])

    def multi_tensor_l2norm(chunk, noop, lists, per_tensor=False):
        f = [int(noop.item())]
        n, p = O.multi_tensor_l2norm(chunk, f, lists, bool(per_tensor))
        noop.fill_(f[0])
        return torch.tensor([float(n)]), (torch.from_numpy(p) if p is not None else torch.empty(0))

    amp_c.multi_tensor_scale, amp_c.multi_tensor_l2norm = multi_tensor_scale, multi_tensor_l2norm
    lamb = types.ModuleType("distributed_lamb_cuda")

    def compute_update_term(chunk, noop, lists, b1, b2, b3, bc, step, eps, mode, decay, gscale, gnorm, max_norm, update_sumsq=None):
        s = DL.compute_update_term([int(noop.item())], lists, b1.tolist(), b2.tolist(), bc.tolist(), int(step), eps.tolist(), mode,
                                   decay.tolist(), float(gscale), float(gnorm), max_norm)
        if s is not None and update_sumsq is not None:
            update_sumsq[:len(s)] = torch.from_numpy(s)

    def update_weights(chunk, noop, lists, pn, un, off, lr, decay, gnorm, nvlamb):
        DL.update_weights([int(noop.item())], lists, pn.tolist(), un.tolist(), off.tolist(), float(lr), decay.tolist(), nvlamb)

    lamb.multi_tensor_lamb_compute_update_term, lamb.multi_tensor_lamb_update_weights = compute_update_term, update_weights
    fac = types.ModuleType("fused_adam_cuda")
    fac.maybe_cast_mt = lambda chunk, flag, lists: O.maybe_cast_mt([int(flag.item())], lists)
    dac = types.ModuleType("distributed_adam_cuda")
    sys.modules.update({"amp_C": amp_c, "distributed_lamb_cuda": lamb, "fused_adam_cuda": fac, "distributed_adam_cuda": dac})


def _ref_lamb_step(masters, m, v, grads, step, lr, b1, b2, eps, wd, max_norm, nvlamb):
    """Unsharded LAMB in float64 on the averaged, unscaled gradients (csrc/multi_tensor_lamb.cu:110-142,255-263)."""
    gn = torch.sqrt(sum((g.double() ** 2).sum() for g in grads))
    clip = min(1.0, max_norm / (float(gn) + 1e-6)) if max_norm > 0 else 1.0
    for i, (p, g) in enumerate(zip(masters, grads)):
        sg = g.double() * clip
        m[i] = b1 * m[i] + (1 - b1) * sg
        v[i] = b2 * v[i] + (1 - b2) * sg * sg
        u = (m[i] / (1 - b1 ** step)) / ((v[i] / (1 - b2 ** step)).sqrt() + eps) + wd * p
        ratio = lr
        if nvlamb or wd != 0:
            pn, un = p.norm(), u.norm()
            ratio = lr * float(pn / un) if (pn != 0 and un != 0) else lr
        p -= ratio * u


def _worker(rank, world, port, backend, out, group_size=0):
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
            _install_cpu_doubles()
        from apex.contrib.optimizers import DistributedFusedLAMB
        g0 = torch.Generator().manual_seed(11)
        init = [torch.randn(s, generator=g0) * 0.1 for s in SHAPES]
        params = [torch.nn.Parameter(t.clone().to(dev)) for t in init]       # fp32 model params: exact comparison
        lr, b1, b2, eps, wd, max_norm = 1e-2, 0.9, 0.99, 1e-6, 0.01, 1.0
        opt = DistributedFusedLAMB(params, lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd, max_grad_norm=max_norm,
                                   use_nvlamb=False, overlap_reductions=False, dwu_group_size=group_size)
        opt.set_global_scale(4.0)
        if group_size:
            assert opt._world_size == group_size and opt._grad_world == world
        masters = [t.double() for t in init]
        m = [torch.zeros_like(t) for t in masters]
        v = [torch.zeros_like(t) for t in masters]
        good_steps = 0
        for it in range(1, 6):
            per_rank = []
            for r in range(world):
                gr = torch.Generator().manual_seed(100 * it + r)
                per_rank.append([torch.randn(s, generator=gr) * 4.0 for s in SHAPES])
            if it == 3:     # one rank produces an inf: every rank must skip, nothing may change, step must not advance
                per_rank[world - 1][2][0] = float("inf")
            for p, g in zip(params, per_rank[rank]):
                p.grad = g.clone().to(dev)
            before = [p.detach().clone() for p in params]
            opt.complete_reductions()
            opt.step()
            if it == 3:
                assert int(opt._overflow_buf) == 1
                for a, b in zip(before, params):
                    assert torch.equal(a, b.detach())
                assert int(opt._step_t) == good_steps
                continue
            good_steps += 1
            assert int(opt._overflow_buf) == 0 and int(opt._step_t) == good_steps
            avg = [sum(gs[i] / world for gs in per_rank) / 4.0 for i in range(len(SHAPES))]
            want_gn = torch.sqrt(sum((a.double() ** 2).sum() for a in avg)) * 4.0
            assert abs(float(opt.L2_grad_norm) - float(want_gn)) <= 1e-5 * float(want_gn)
            _ref_lamb_step(masters, m, v, avg, good_steps, lr, b1, b2, eps, wd, max_norm, False)
            for p, w in zip(params, masters):
                assert torch.allclose(p.detach().cpu().double(), w, rtol=2e-5, atol=2e-6), (it, rank)
        sd = opt.state_dict()
        assert int(sd["step"]) == 4 and sd["fp32_p"].numel() == opt._layout["shard"]
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _worker_e5m2(rank, world, port, backend, out):
    """e5m2_allgather=True next to a plain optimizer: an overflowing first step gathers the packed initial
    parameters, the first clean step leaves bit-identical fp32 shards, and the model parameters are the e5m2
    rounding of the masters (reference distributed_fused_lamb.py:203-206,850-861)."""
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
            _install_cpu_doubles()
        from apex.contrib.optimizers import DistributedFusedLAMB
        from oracle import mt_oracle as O
        g0 = torch.Generator().manual_seed(11)
        init = [torch.randn(s, generator=g0) * 0.1 for s in SHAPES]
        pa = [torch.nn.Parameter(t.clone().to(dev).half()) for t in init]
        pb = [torch.nn.Parameter(t.clone().to(dev).half()) for t in init]
        kw = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-6, weight_decay=0.01, max_grad_norm=1.0, overlap_reductions=False)
        oa, ob = DistributedFusedLAMB(pa, e5m2_allgather=True, **kw), DistributedFusedLAMB(pb, **kw)
        for it in range(2):
            gr = torch.Generator().manual_seed(100 * it + rank)
            grads = [(torch.randn(s, generator=gr) * 4.0).half() for s in SHAPES]
            if it == 0 and rank == 1:
                grads[2][0] = float("inf")
            for a, b, g in zip(pa, pb, grads):
                a.grad, b.grad = g.clone().to(dev), g.clone().to(dev)
            for o in (oa, ob):
                o.set_global_scale(4.0)
                o.complete_reductions()
                o.step()
            assert oa._new_params.dtype == torch.uint8 and int(oa._overflow_buf) == int(ob._overflow_buf) == (1 if it == 0 else 0)
            shard = oa._layout["shard"]
            mine = oa._new_params[rank * shard:(rank + 1) * shard].cpu().numpy()
            assert (mine == O.pack_e5m2(oa._fp32_p.cpu().numpy())).all()
            if it == 0:      # skipped step: model parameters untouched, gathered bytes = packed initial parameters
                for a, t in zip(pa, init):
                    assert torch.equal(a.detach().cpu(), t.half())
            else:
                assert torch.equal(oa._fp32_p, ob._fp32_p) and torch.equal(oa._fp32_m, ob._fp32_m)
                for a, b in zip(pa, pb):
                    h = a.detach().view(torch.int16)
                    assert bool(((h & 0xFF) == 0).all())
                    d = (a.detach().float() - b.detach().float()).abs()
                    assert bool((d <= 0.13 * b.detach().float().abs() + 2e-5).all())
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _run(backend, port, worker=None, world=2, extra=()):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker or _worker, args=(r, world, port, backend, q) + tuple(extra)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def test_sharded_lamb_host_logic_world2_gloo():
    _run("gloo", 29750)


def test_sharded_lamb_e5m2_allgather_world2_gloo():
    _run("gloo", 29752, _worker_e5m2)


def test_sharded_lamb_rank_groups_world4_gloo():
    """dwu_group_size=2 on 4 ranks: shards inside a group, gradients all-reduced across groups
    (reference distributed_fused_lamb.py:127-130,148-168)."""
    _run("gloo", 29754, world=4, extra=(2,))


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_lamb_e5m2_allgather_world2_nccl():
    _run("nccl", 29753, _worker_e5m2)


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_lamb_matches_unsharded_world2_nccl():
    _run("nccl", 29751)


@pytest.mark.gpu
def test_sharded_lamb_world1_nccl_one_gpu():
    """1-rank NCCL job on one GPU: the class, its collectives and both sharded-LAMB kernels against float64 LAMB."""
    _run("nccl", 29755, world=1)


@pytest.mark.gpu
def test_dist_lamb_kernels_match_oracle():
    _paths()
    import distributed_lamb_cuda as ours
    from oracle import dist_lamb_oracle as DL
    from util import assert_close_rel
    sizes = [1, 3, 555, 65537, 33333, 8192]
    T = len(sizes)
    g = torch.Generator(device="cuda").manual_seed(70)
    mk = lambda dt, s=1.0, pos=False: [((torch.randn(n, generator=g, device="cuda") * s).abs() if pos else torch.randn(n, generator=g, device="cuda") * s).to(dt) for n in sizes]
    for gdt in (torch.float16, torch.float32, torch.bfloat16):
        for mode in (0, 1):
            lists = [mk(gdt, 64.0), mk(torch.float32, 0.1), mk(torch.float32, 0.01), mk(torch.float32, 0.001, True), mk(torch.float32)]
            c = [[t.cpu().clone() for t in l] for l in lists]
            b1 = torch.linspace(0.85, 0.95, T, device="cuda"); b2 = torch.linspace(0.98, 0.999, T, device="cuda")
            bc = torch.tensor([1, 0] * (T // 2), device="cuda", dtype=torch.int32)
            eps = torch.full((T,), 1e-6, device="cuda"); wd = torch.tensor([0.0, 0.01] * (T // 2), device="cuda")
            step = torch.tensor([3], device="cuda", dtype=torch.int32)
            gscale, gnorm = torch.tensor([64.0], device="cuda"), torch.tensor([640.0], device="cuda")
            sumsq = torch.zeros(T, device="cuda")
            flag = torch.zeros(1, dtype=torch.int32, device="cuda")
            ours.multi_tensor_lamb_compute_update_term(65536, flag, lists, b1, b2, 1 - b1, bc, step, eps, mode, wd, gscale, gnorm, 1.0, sumsq)
            want = DL.compute_update_term([0], c, b1.tolist(), b2.tolist(), bc.tolist(), 3, eps.tolist(), mode, wd.tolist(), 64.0, 640.0, 1.0)
            for k in (2, 3, 4):
                for a, b in zip(lists[k], c[k]):
                    assert_close_rel(a, b, rtol=2e-6, scale="self", what="stage1 list %d" % k)
            assert_close_rel(sumsq, torch.from_numpy(want), rtol=1e-5, what="per-fragment sum of squares of u")
            # stage 2, fp16 and e5m2 copies
            for cdt in (torch.float16, torch.uint8, torch.float32):
                u, p = [t.clone() for t in lists[4]], [t.clone() for t in lists[1]]
                pc = [torch.zeros(n, dtype=cdt, device="cuda") for n in sizes]
                cu, cp = [t.cpu().clone() for t in u], [t.cpu().clone() for t in p]
                cpc = [torch.zeros(n, dtype=torch.float16 if cdt == torch.uint8 else cdt) for n in sizes]
                pn = torch.linspace(1.0, 2.0, T, device="cuda"); un = torch.linspace(0.5, 3.0, T, device="cuda")
                off = torch.tensor(list(reversed(range(T))), device="cuda", dtype=torch.int64)
                lrt = torch.tensor([1e-2], device="cuda")
                ours.multi_tensor_lamb_update_weights(65536, flag, [u, p, pc], pn, un, off, lrt, wd, gnorm, False)
                DL.update_weights([0], [cu, cp, cpc], pn.tolist(), un.tolist(), off.tolist(), 1e-2, wd.tolist(), False)
                for a, b in zip(p, cp):
                    assert_close_rel(a, b, rtol=1e-6, scale="self", what="stage2 p")
                if cdt != torch.uint8:
                    for a, b in zip(pc, cpc):
                        assert_close_rel(a, b, rtol=1e-3 if cdt == torch.float16 else 1e-6, scale="self", what="stage2 p_copy")
                else:   # the bytes are the (reference-pinned) e5m2 packing of the kernel's own fp32 result
                    from oracle import mt_oracle as O
                    import numpy as np
                    for a, b in zip(pc, p):
                        assert np.array_equal(a.cpu().numpy(), O.pack_e5m2(b.cpu().numpy()))
            # gating: flag == 1 -> nothing moves
            keep = [[t.clone() for t in l] for l in lists]
            one = torch.ones(1, dtype=torch.int32, device="cuda")
            ours.multi_tensor_lamb_compute_update_term(65536, one, lists, b1, b2, 1 - b1, bc, step, eps, mode, wd, gscale, gnorm, 1.0, sumsq)
            for l0, l1 in zip(keep, lists):
                assert all(torch.equal(a, b) for a, b in zip(l0, l1))


@pytest.mark.gpu
def test_dist_lamb_kernels_match_reference_binary():
    """ops_dist_lamb.cu and oracle/dist_lamb_oracle.py against the reference's own
    multi_tensor_distopt_lamb_kernel.cu compiled unmodified for sm_100a (oracle/_ref/distributed_lamb_cuda_ref.so,
    oracle/build_ref.sh): both kernels, fp32 and fp16 gradients (the reference dispatches float / half only,
    multi_tensor_distopt_lamb_kernel.cu:446-449), both decay modes, clipping on and off, fp32 / fp16 / e5m2-byte
    parameter copies, and the noop gate.  Tolerances: v (a sum of non-negative terms) pure element-wise 2e-6,
    m / u / p element-wise 1e-6 plus 1e-6 of the tensor's largest magnitude (cancellation), bytes bit-exact."""
    _paths()
    import numpy as np
    import distributed_lamb_cuda as ours
    from oracle import dist_lamb_oracle as DL
    from util import assert_close_rel, assert_rel_pure, load_ref, rel_err_stats
    refm = load_ref("distributed_lamb_cuda_ref")
    if refm is None:
        pytest.skip("oracle/_ref/distributed_lamb_cuda_ref.so not built")
    sizes = [1, 3, 555, 65537, 33333, 8192, 262144, 4]
    T = len(sizes)
    g = torch.Generator(device="cuda").manual_seed(71)

    def mk(dt, s=1.0, pos=False):
        out = []
        for n in sizes:
            x = torch.randn(n, generator=g, device="cuda") * s
            out.append((x.abs() if pos else x).to(dt))
        return out

    worst = {}
    for gdt in (torch.float16, torch.float32):
        for mode in (0, 1):
            for max_norm in (1.0, 0.0):
                lists = [mk(gdt, 64.0), mk(torch.float32, 0.1), mk(torch.float32, 0.01), mk(torch.float32, 0.001, True),
                         mk(torch.float32)]
                rl = [[t.clone() for t in l] for l in lists]
                cl = [[t.cpu().clone() for t in l] for l in lists]
                b1 = torch.linspace(0.85, 0.95, T, device="cuda"); b2 = torch.linspace(0.98, 0.999, T, device="cuda")
                bc = torch.tensor([1, 0] * (T // 2), device="cuda", dtype=torch.int32)
                eps = torch.full((T,), 1e-6, device="cuda"); wd = torch.tensor([0.0, 0.01] * (T // 2), device="cuda")
                step = torch.tensor([3], device="cuda", dtype=torch.int32)
                gscale, gnorm = torch.tensor([64.0], device="cuda"), torch.tensor([640.0], device="cuda")
                flag, rflag = (torch.zeros(1, dtype=torch.int32, device="cuda") for _ in range(2))
                ours.multi_tensor_lamb_compute_update_term(65536, flag, lists, b1, b2, 1 - b1, bc, step, eps, mode, wd, gscale,
                                                           gnorm, max_norm)
                refm.multi_tensor_lamb_compute_update_term(65536, rflag, rl, b1, b2, 1 - b1, bc, step, eps, mode, wd, gscale,
                                                           gnorm, max_norm)
                DL.compute_update_term([0], cl, b1.tolist(), b2.tolist(), bc.tolist(), 3, eps.tolist(), mode, wd.tolist(), 64.0,
                                       640.0, max_norm)
                assert int(flag) == int(rflag) == 0
                for a, b in zip(lists[0], rl[0]):
                    assert torch.equal(a, b)                 # gradients are inputs only in both
                assert_rel_pure(lists[3], rl[3], 2e-6, "stage1 v vs reference binary")
                assert_rel_pure([t.cuda() for t in cl[3]], rl[3], 2e-6, "oracle v vs reference binary")
                for k, name in ((2, "m"), (4, "u")):
                    for a, b, c in zip(lists[k], rl[k], cl[k]):
                        assert_close_rel(a, b, rtol=1e-6, scale="self", what="stage1 %s vs reference binary" % name)
                        assert_close_rel(c.cuda(), b, rtol=2e-6, scale="self", what="oracle %s vs reference binary" % name)
                for k, name in ((2, "m"), (3, "v"), (4, "u")):
                    mx, p999, _ = rel_err_stats(lists[k], rl[k], floor=1e-12)
                    worst[name] = max(worst.get(name, 0.0), p999)
                # stage 2 on the reference's update, so that both sides start from identical inputs
                for cdt in (torch.float16, torch.uint8, torch.float32):
                    for nvlamb in (False, True):
                        u = [t.clone() for t in rl[4]]
                        p, rp = [t.clone() for t in rl[1]], [t.clone() for t in rl[1]]
                        pc, rpc = ([torch.zeros(n, dtype=cdt, device="cuda") for n in sizes] for _ in range(2))
                        pn = torch.linspace(1.0, 2.0, T, device="cuda"); un = torch.linspace(0.5, 3.0, T, device="cuda")
                        un[2] = 0.0                                                       # zero norm: plain lr
                        off = torch.tensor(list(reversed(range(T))), device="cuda", dtype=torch.int64)
                        lrt = torch.tensor([1e-2], device="cuda")
                        ours.multi_tensor_lamb_update_weights(65536, flag, [u, p, pc], pn, un, off, lrt, wd, gnorm, nvlamb)
                        refm.multi_tensor_lamb_update_weights(65536, rflag, [[t.clone() for t in u], rp, rpc], pn, un, off, lrt,
                                                              wd, gnorm, nvlamb)
                        for a, b in zip(p, rp):
                            assert_close_rel(a, b, rtol=1e-6, scale="self", what="stage2 p vs reference binary")
                        for a, b, pa_ in zip(pc, rpc, p):
                            if cdt == torch.uint8:
                                # bytes: equal wherever the two fp32 results are (they may differ in the last bit of p)
                                assert np.array_equal(a.cpu().numpy(), O_pack(pa_))
                            elif cdt == torch.float16:
                                assert_close_rel(a, b, rtol=1e-3, scale="self", what="stage2 fp16 copy vs reference binary")
                            else:
                                assert_close_rel(a, b, rtol=1e-6, scale="self", what="stage2 fp32 copy vs reference binary")
                        if cdt == torch.uint8:
                            # and the reference's own bytes are the same packing of ITS fp32 result
                            for b, pr in zip(rpc, rp):
                                assert np.array_equal(b.cpu().numpy(), O_pack(pr))
                # noop gate: flag == 1 -> neither implementation moves anything
                keep = [[t.clone() for t in l] for l in lists]
                one = torch.ones(1, dtype=torch.int32, device="cuda")
                ours.multi_tensor_lamb_compute_update_term(65536, one, lists, b1, b2, 1 - b1, bc, step, eps, mode, wd, gscale, gnorm, max_norm)
                refm.multi_tensor_lamb_compute_update_term(65536, one, rl, b1, b2, 1 - b1, bc, step, eps, mode, wd, gscale, gnorm, max_norm)
                for l0, l1 in zip(keep, lists):
                    assert all(torch.equal(a, b) for a, b in zip(l0, l1))
    print("dist LAMB vs reference binary, 99.9th-percentile relative error: %s" % worst)


def O_pack(t):
    from oracle import mt_oracle as O
    return O.pack_e5m2(t.detach().cpu().numpy())
