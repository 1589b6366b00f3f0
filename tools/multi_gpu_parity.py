"""Multi-GPU correctness checks that run INSIDE an existing torch.distributed (NCCL) job -- bench.py's N > 1 arm
calls them before its timed loop, so every SCALE line carries a `parity` block measured on the same ranks,
buffers and kernels that are then timed.  (The pytest versions of these checks, tests/test_dist_adam_gpu.py,
tests/test_dist_lamb.py, tests/test_ddp.py, spawn their own 2-rank jobs and are skipped on a 1-GPU box.)

What the reference checks for this path is tests/L0/run_optimizers/test_dist_adam.py:119-173: N-rank
DistributedFusedAdam against the unsharded optimizer, allclose on the weights.  Here:

  sharded_adam_parity   one step of DistributedFusedAdam (NCCL pipeline or the fused multimem kernel) on the FULL
                        bench workload, compared on a deterministic sample of fragments of every rank's shard (the
                        fragments cut by the shard boundaries at both ends, and one in the middle) with the UNSHARDED
                        kernel b200mt_dist_adam (pinned to the reference's own kernel on one GPU by
                        tests/test_parity_gpu.py) applied to the same reduced gradient; plus the reduced gradient
                        itself against an fp32 all-reduce, and the gathered parameters on EVERY rank.
  sharded_lamb_parity   DistributedFusedLAMB over the job's ranks against an unsharded float64 LAMB, including a step
                        with an inf on one rank that every rank must skip.
  ddp_parity            apex.parallel.DistributedDataParallel bucket path: every rank ends with the mean gradient.

Nothing here imports oracle/: the comparisons are against this library's single-GPU kernels and float64 torch.
"""
import math
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def _ord16(t):
    """16-bit float bit patterns -> integers whose difference counts representable values (ulps)."""
    i = t.contiguous().view(torch.int16).to(torch.int32)
    return torch.where(i < 0, -(i & 0x7FFF), i)


def ulp_diff16(a, b):
    """max over elements of the distance, in representable values, between two bf16/fp16 tensors (uint8: byte distance)."""
    if a.numel() == 0:
        return 0
    if a.dtype == torch.uint8:
        return int((a.to(torch.int32) - b.to(torch.int32)).abs().max())
    return int((_ord16(a) - _ord16(b)).abs().max())


def max_rel(a, b):
    """max |a-b| / |b| over elements with b != 0 (fp64); elements with b == 0 must match exactly."""
    a, b = a.double().reshape(-1), b.double().reshape(-1)
    if a.numel() == 0:
        return 0.0
    nz = b != 0
    worst = float(((a[nz] - b[nz]).abs() / b[nz].abs()).max()) if bool(nz.any()) else 0.0
    if bool((~nz).any()) and float(a[~nz].abs().max()) != 0.0:
        return float("inf")
    return worst


def _pick_samples(frags, K, fused, sample_elems):
    """(chunk k, fragment j, first element a, end element b) -- a, b multiples of 64 or the fragment's end."""
    out, seen = [], set()

    def add(k, j, head):
        n = frags[k][j][3]
        if fused:
            n = (n + 7) // 8 * 8                     # the fused kernel's fragments extend into the alignment padding
        m = min(n, sample_elems)
        a, b = (0, m) if head else ((n - m) // 64 * 64, n)
        if (k, j, a, b) not in seen:
            seen.add((k, j, a, b))
            out.append((k, j, a, b))

    for k in sorted({0, K // 2, K - 1}):
        if not frags[k]:
            continue
        add(k, 0, True)                              # a parameter cut by the start of this rank's shard
        add(k, len(frags[k]) - 1, False)             # ... and by its end
        if len(frags[k]) > 2:
            add(k, len(frags[k]) // 2, True)
    return out


@torch.no_grad()
def sharded_adam_parity(opt, sample_elems=16 << 20):
    """Runs ONE optimizer step of `opt` (a DistributedFusedAdam whose flat gradient buffer is already filled) and
    returns (verdict dict, ok).  Collective: every rank of the job must call it."""
    import _b200mt as mt
    import distributed_adam_cuda
    from apex.contrib.optimizers.distributed_fused_adam import shard_fragments

    opt._init_everything()
    L, K, dev = opt._layout, opt._num_chunks_total, opt._device
    shard, chunk, W, r = L["shard"], L["chunk"], opt._world_size, opt._rank
    fused, low = opt._fused, opt._low
    numels = [p.numel() for p in opt._model_params]
    frags = shard_fragments(L, numels, r)
    mine = _pick_samples(frags, K, fused, sample_elems)
    # (flat position, local position, length, hyper-parameter row) of every sample of this rank
    rows = []
    for (k, j, a, b) in mine:
        soff = frags[k][j][2]
        rows.append({"flat": k * chunk + r * shard + soff + a, "local": k * shard + soff + a, "n": b - a, "k": k, "j": j})
    p0 = [opt._fp32_p[x["local"]:x["local"] + x["n"]].clone() for x in rows]
    m0 = [opt._fp32_m[x["local"]:x["local"] + x["n"]].clone() for x in rows]
    v0 = [opt._fp32_v[x["local"]:x["local"] + x["n"]].clone() for x in rows]
    hyper = [torch.stack([opt._chunk_hyper[x["k"]][h][x["j"]] for x in rows]) for h in range(5)]

    # ---- the gradient every sample's owner should see: fp32 sum over the ranks (and the sum of magnitudes)
    everyone = [None] * W
    dist.all_gather_object(everyone, [(x["flat"], x["n"]) for x in rows], group=opt._pg)
    gsum, gabs = [], []
    for owner in range(W):
        for (pos, n) in everyone[owner]:
            c = opt._flat_grads[pos:pos + n].float()
            s, a = c.clone(), c.abs()
            dist.all_reduce(s, group=opt._pg)
            dist.all_reduce(a, group=opt._pg)
            if owner == r:
                gsum.append(s); gabs.append(a)
            del c, s, a

    # ---- the step under test
    opt.complete_reductions()
    opt.step()
    sync = (lambda: torch.cuda.synchronize(dev)) if opt._cuda else (lambda: None)
    sync()
    step_no, lr = opt._step, opt.param_groups[0]["lr"]
    scale = opt._combined_scale()

    # ---- the reduced gradient the step kernel consumed
    if fused and getattr(opt, "_fused_form", None) == "p2p":
        # two ranks over peer memory: the kernel's load IS fp32(own) + fp32(peer) rounded once to the gradient type --
        # bit for bit the fp32 all-reduce above, rounded
        gred = [s_.to(low) for s_ in gsum]
    elif fused:
        # never materialised by the fused kernel: repeat exactly its load (multimem.ld_reduce through the same
        # multicast addresses) with the stand-alone reduce-scatter kernel; the gradients are untouched by the step
        esz = opt._flat_grads.element_size()
        gred = [torch.empty(x["n"], dtype=low, device=dev) for x in rows]
        opt._hdl_g.barrier(channel=0)
        if rows:
            src = [mt.RawBuffer(opt._hdl_g.multicast_ptr + x["flat"] * esz, x["n"], low, dev) for x in rows]
            Lrs = mt.pack([src, gred], 2)
            mt.check(mt.lib.b200mt_nvls_reduce_scatter(Lrs.ref(), 65536, torch.cuda.current_stream(dev).cuda_stream))
        opt._hdl_g.barrier(channel=1)
        sync()
    else:
        gred = [opt._fp16_g[x["local"]:x["local"] + x["n"]].clone() for x in rows]

    # ---- unsharded kernel on (state before the step, that reduced gradient)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    copy_dt = torch.uint8 if opt._e5m2 else low
    pc_ref = [torch.zeros(x["n"], dtype=copy_dt, device=dev) for x in rows]
    pc_ideal = [torch.zeros(x["n"], dtype=copy_dt, device=dev) for x in rows]
    res = {"master_p_m_v_max_rel": 0.0, "copy_vs_unsharded_max_ulp": 0, "reduced_grad_max_err_over_sum_abs": 0.0,
           "reduced_grad_vs_rn_fp32_sum_max_ulp": 0, "elements_checked": 0}
    pa, ma, va = ([t.clone() for t in l] for l in (p0, m0, v0))
    gideal = [s_.to(low) for s_ in gsum]
    if rows:
        distributed_adam_cuda.multi_tensor_fused_adam(65536, flag, [pa, ma, va, gred, pc_ref], hyper[0], hyper[1], hyper[2],
                                                      hyper[3], hyper[4], lr, scale, step_no, opt.eps_mode)
        # the same from the "ideal" gradient: ONE rounding of the fp32 sum (what an fp32 all-reduce would hand over);
        # reported, not bounded: where the 16-bit reduction rounded differently the update differs by design
        distributed_adam_cuda.multi_tensor_fused_adam(65536, flag, [p0, m0, v0, gideal, pc_ideal], hyper[0], hyper[1], hyper[2],
                                                      hyper[3], hyper[4], lr, scale, step_no, opt.eps_mode)
        sync()
    # pa/ma/va: unsharded result from the same reduced gradient; p0/m0/v0: from the ideal gradient
    for i, x in enumerate(rows):
        sl = slice(x["local"], x["local"] + x["n"])
        for got, want in ((opt._fp32_p[sl], pa[i]), (opt._fp32_m[sl], ma[i]), (opt._fp32_v[sl], va[i])):
            res["master_p_m_v_max_rel"] = max(res["master_p_m_v_max_rel"], max_rel(got, want))
        if not fused:
            res["copy_vs_unsharded_max_ulp"] = max(res["copy_vs_unsharded_max_ulp"], ulp_diff16(opt._fp16_p[sl], pc_ref[i]))
        # informational (not bounded): against ONE rounding of the fp32 sum.  ulps only where the sum is not a cancellation
        # (|sum| >= sum|g_i| / 4: elsewhere a 16-bit ulp of the tiny result says nothing), and how many elements of the
        # low-precision copy come out identical to that path
        solid = gsum[i].abs() >= 0.25 * gabs[i]
        if bool(solid.any()):
            res["reduced_grad_vs_rn_fp32_sum_max_ulp"] = max(res["reduced_grad_vs_rn_fp32_sum_max_ulp"],
                                                             ulp_diff16(gred[i][solid], gideal[i][solid]))
        same = (opt._new_params[x["flat"]:x["flat"] + x["n"]] == pc_ideal[i]).sum()
        res["copy_equal_to_fp32_allreduce_path"] = res.get("copy_equal_to_fp32_allreduce_path", 0) + int(same)
        err = (gred[i].float() - gsum[i]).abs() / gabs[i].clamp_min(1e-30)
        res["reduced_grad_max_err_over_sum_abs"] = max(res["reduced_grad_max_err_over_sum_abs"], float(err.max()))
        res["elements_checked"] += x["n"]
    del gsum, gabs

    # ---- the gathered / broadcast parameters, on EVERY rank, against the owner's unsharded result
    for owner in range(W):
        for i, (pos, n) in enumerate(everyone[owner]):
            want = pc_ref[i] if owner == r else torch.empty(n, dtype=copy_dt, device=dev)
            dist.broadcast(want, src=dist.get_global_rank(opt._pg, owner), group=opt._pg)
            res["copy_vs_unsharded_max_ulp"] = max(res["copy_vs_unsharded_max_ulp"], ulp_diff16(opt._new_params[pos:pos + n], want))
    t = torch.tensor([res["master_p_m_v_max_rel"], float(res["copy_vs_unsharded_max_ulp"]), res["reduced_grad_max_err_over_sum_abs"],
                      float(res["reduced_grad_vs_rn_fp32_sum_max_ulp"])], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=opt._pg)
    n = torch.tensor([float(res["elements_checked"]), float(res.get("copy_equal_to_fp32_allreduce_path", 0))], dtype=torch.float64,
                     device=dev)
    dist.all_reduce(n, group=opt._pg)
    res = {"master_p_m_v_max_rel": float(t[0]), "copy_vs_unsharded_max_ulp": int(t[1]),
           "reduced_grad_max_err_over_sum_abs": float(t[2]), "reduced_grad_vs_rn_fp32_sum_max_ulp": int(t[3]),
           "copy_fraction_identical_to_an_fp32_allreduce_path": float(n[1]) / max(1.0, float(n[0])),
           "elements_checked": int(n[0]), "samples_per_rank": len(rows), "step": step_no}
    # bounds: identical inputs through the sharded and the unsharded launch of the same rule -> 1e-6 on the fp32
    # state, one 16-bit ulp on the copy.  A 16-bit sum of W terms is rounded up to W-1 times on the way (NCCL ring: half
    # an ulp of a partial sum each, i.e. <= 2^-9 of sum|g_i| for bf16) and the switch's final rounding may sit one ulp
    # (2^-8 of the sum) off round-to-nearest: (W + 1) * 2^-8 of sum|g_i| covers both with a factor two to spare
    bound_g = (W + 1) * 2.0 ** -8
    res["bounds"] = {"master_p_m_v_max_rel": 1e-6, "copy_max_ulp": 1, "reduced_grad_max_err_over_sum_abs": bound_g}
    ok = (res["master_p_m_v_max_rel"] <= 1e-6 and res["copy_vs_unsharded_max_ulp"] <= 1
          and res["reduced_grad_max_err_over_sum_abs"] <= bound_g and res["elements_checked"] > 0)
    return res, ok


# ------------------------------------------------------------------------------------------------------------------
LAMB_SHAPES = [(1024, 1024), (1024,), (4096, 1024), (4096,), (1000, 3), (1,), (30522, 64), (5,)]


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


def sharded_lamb_parity(dev, rank, world):
    """DistributedFusedLAMB (fp32 model parameters: exact comparison) over the job's ranks vs float64 LAMB; iteration 3
    carries an inf on the last rank.  Returns (dict, ok)."""
    from apex.contrib.optimizers import DistributedFusedLAMB
    g0 = torch.Generator().manual_seed(11)
    init = [torch.randn(s, generator=g0) * 0.1 for s in LAMB_SHAPES]
    params = [torch.nn.Parameter(t.clone().to(dev)) for t in init]
    lr, b1, b2, eps, wd, max_norm = 1e-2, 0.9, 0.99, 1e-6, 0.01, 1.0
    opt = DistributedFusedLAMB(params, lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd, max_grad_norm=max_norm,
                               use_nvlamb=False, overlap_reductions=False)
    opt.set_global_scale(4.0)
    masters = [t.double().to(dev) for t in init]
    m = [torch.zeros_like(t) for t in masters]
    v = [torch.zeros_like(t) for t in masters]
    good, worst, worst_norm, skipped_ok = 0, 0.0, 0.0, False
    for it in range(1, 5):
        per_rank = []
        for rr in range(world):
            gr = torch.Generator().manual_seed(100 * it + rr)
            per_rank.append([torch.randn(s, generator=gr) * 4.0 for s in LAMB_SHAPES])
        if it == 3:
            per_rank[world - 1][2][0] = float("inf")
        for p, g in zip(params, per_rank[rank]):
            p.grad = g.clone().to(dev)
        before = [p.detach().clone() for p in params]
        opt.complete_reductions()
        opt.step()
        if it == 3:
            skipped_ok = (int(opt._overflow_buf) == 1 and int(opt._step_t) == good
                          and all(torch.equal(a, b.detach()) for a, b in zip(before, params)))
            continue
        good += 1
        if int(opt._overflow_buf) != 0 or int(opt._step_t) != good:
            return {"error": "flag %d / step %d after clean iteration %d" % (int(opt._overflow_buf), int(opt._step_t), it)}, False
        avg = [(sum(gs[i] for gs in per_rank) / world / 4.0).to(dev) for i in range(len(LAMB_SHAPES))]
        want_gn = float(torch.sqrt(sum((a.double() ** 2).sum() for a in avg))) * 4.0
        worst_norm = max(worst_norm, abs(float(opt.L2_grad_norm) - want_gn) / want_gn)
        _ref_lamb_step(masters, m, v, avg, good, lr, b1, b2, eps, wd, max_norm, False)
        for p, w in zip(params, masters):
            d = (p.detach().double() - w).abs() / (w.abs() + 0.1)     # |p| ~ 0.1: error relative to the parameter scale
            worst = max(worst, float(d.max()))
    t = torch.tensor([worst, worst_norm, 0.0 if skipped_ok else 1.0], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = {"params": sum(math.prod(s) for s in LAMB_SHAPES), "clean_steps": good, "param_max_err_rel_to_scale": float(t[0]),
           "grad_norm_max_rel": float(t[1]), "overflow_step_skipped_on_every_rank": float(t[2]) == 0.0,
           "bounds": {"param_max_err_rel_to_scale": 2e-5, "grad_norm_max_rel": 1e-5}}
    ok = res["param_max_err_rel_to_scale"] <= 2e-5 and res["grad_norm_max_rel"] <= 1e-5 and res["overflow_step_skipped_on_every_rank"]
    return res, ok


# ------------------------------------------------------------------------------------------------------------------
class _Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(256, 512)
        self.b = torch.nn.Linear(512, 512)
        self.c = torch.nn.Linear(512, 64)

    def forward(self, x):
        return self.c(torch.relu(self.b(torch.relu(self.a(x)))))


def _local_grads(model, x):
    for p in model.parameters():
        p.grad = None
    model(x).pow(2).mean().backward()
    return [p.grad.detach().clone() for p in model.parameters()]


def ddp_parity(dev, rank, world):
    """apex.parallel.DistributedDataParallel: after backward every rank holds the mean of the per-rank gradients
    (overlapped buckets, delayed all-reduce, predivide).  Returns (dict, ok)."""
    import copy
    from apex.parallel import DistributedDataParallel as DDP
    torch.manual_seed(100 + rank)
    base = _Net().to(dev)
    worst, cases = 0.0, 0
    for kw in (dict(message_size=100000), dict(delay_allreduce=True), dict(message_size=1, gradient_predivide_factor=2.0)):
        model = copy.deepcopy(base)
        ddp = DDP(model, **kw)
        ref_model = copy.deepcopy(model)
        for it in range(3):
            xs = [torch.randn(32, 256, generator=torch.Generator().manual_seed(1000 * it + rr)).to(dev) for rr in range(world)]
            want = [_local_grads(ref_model, x) for x in xs]
            for p in model.parameters():
                p.grad = None
            ddp(xs[rank]).pow(2).mean().backward()
            torch.cuda.synchronize(dev)
            for i, p in enumerate(model.parameters()):
                mean = sum(w[i] for w in want) / world
                scale = float(mean.abs().max()) + 1e-12
                worst = max(worst, float((p.grad - mean).abs().max()) / scale)
            cases += 1
    t = torch.tensor([worst], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = {"cases": cases, "grad_max_err_rel_to_largest": float(t[0]), "bounds": {"grad_max_err_rel_to_largest": 1e-5}}
    return res, res["grad_max_err_rel_to_largest"] <= 1e-5
