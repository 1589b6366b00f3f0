"""N > 1 arm of bench.py: BASELINE.json configs[3] -- the ZeRO-style sharded optimizer step of
DistributedFusedAdam on the Llama-3-8B shape list (291 tensors, 8,030,261,248 params; bf16 gradients and
model parameters, fp32 master weights and moments sharded over the ranks):

    reduce-scatter(bf16 grads) -> shard-local fused AdamW (one RMW pass, writes the bf16 copy) -> all-gather(bf16 params)

One process per GPU (torch.distributed.run), NCCL over NVLink/NVSwitch.  Strong scaling: the list is fixed,
each rank owns 1/N of it.  Times are CUDA events on each rank, max over ranks.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

import param_shapes

NVLINK_GBS = 770.0     # measured peer-copy bandwidth per direction (B200_PROFILING.md); 900 nominal


def _report_fused(args, opt, ms, launches, sampler, rank, world, local, dev, stream, flat, nparams, shapes, model_name, lib,
                  parity=None, parity_ok=True, nvlink_measured=None):
    from bench import peaks, pinned_near_gpu, gpu_numa_cpus, host_numa_nodes, METRIC, UNIT
    clocks = sampler.stop() if rank == 0 else None
    host = pinned_near_gpu(1 << 27, torch.bfloat16, local)
    numa_known = gpu_numa_cpus(local) is not None or host_numa_nodes() == 1
    host.normal_(0, 1)
    e2e_steps = max(2, min(args.steps, 5))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def e2e_step():
        for off in range(0, flat.numel(), host.numel()):
            n = min(host.numel(), flat.numel() - off)
            flat[off:off + n].copy_(host[:n], non_blocking=True)
        opt.complete_reductions()
        opt.step()
        return float(opt._new_params[0])

    e2e_step()
    torch.cuda.synchronize(); dist.barrier()
    e0.record(stream)
    for _ in range(e2e_steps):
        e2e_step()
    e1.record(stream)
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tl = torch.tensor([float(launches)], device=dev, dtype=torch.float64)
    dist.all_reduce(tl, op=dist.ReduceOp.SUM)
    ms, e2e_ms = t.tolist()
    if rank == 0:
        peak, peak_src = peaks()
        total = opt._layout["total"]
        shard_elems = opt._layout["shard"]
        # per GPU and direction: out = my whole gradient buffer pulled by the switch (2*total) + my shard's
        # parameters sent once (2*shard); in = my shard's reduced gradients (2*shard) + everybody's parameters (2*total)
        pbytes = 1.0 if opt._e5m2 else 2.0     # bytes per parameter in the broadcast
        p2p = getattr(opt, "_fused_form", None) == "p2p"
        if p2p:     # two ranks over peer memory: the peer reads my copy of its shard (2*shard out), I write my shard's parameters to it
            nv_bytes = 2.0 * shard_elems + pbytes * shard_elems
        else:
            nv_bytes = 2.0 * total + pbytes * shard_elems      # the busier (outgoing) direction
        kbytes = (26.0 + pbytes) * shard_elems
        achieved = kbytes / (ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": nparams / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": ("configs[3]: DistributedFusedAdam sharded step on the %s shape list as ONE kernel over peer-mapped memory "
                                    "(two ranks): the peer's bf16 grads added on load, fused AdamW (fp32 master/moments), bf16 params stored "
                                    "here and in the peer" % model_name) if p2p else
                                   ("configs[3]: DistributedFusedAdam sharded step on the %s shape list as ONE kernel over NVSwitch multicast: "
                                    "bf16 grads reduced in the switch on load (multimem.ld_reduce), fused AdamW (fp32 master/moments), "
                                    "bf16 params broadcast by the switch on store (multimem.st)" % model_name),
                       "fused_form": getattr(opt, "_fused_form", None),
                       "tensors": len(shapes), "params": nparams, "flat_elems": total, "parallelism": "zero-dp%d" % world,
                       "e5m2_allgather": bool(opt._e5m2),
                       "l2": "inputs larger than L2 (%.1f GB of optimizer state per rank)" % (12.0 * shard_elems / 1e9)},
            "roofline": {"bound": "nvlink", "achieved": nv_bytes / (ms * 1e-3) / 1e9, "peak": NVLINK_GBS, "unit": "GB/s",
                         "frac": nv_bytes / (ms * 1e-3) / 1e9 / NVLINK_GBS,
                         "traffic": (nvlink_measured or {}).get("tx_bytes_per_step_per_gpu"),
                         "traffic_source": ("NVLink bytes this GPU transmitted per step (NVML link counters around the timed loop), "
                                            "the counterpart of `bytes_per_launch` for an NVLink-bound kernel")
                         if (nvlink_measured or {}).get("tx_bytes_per_step_per_gpu") is not None else
                         ("null: this box's NVML does not expose the NVLink throughput counters (%s); the single-collective "
                          "timings of tools/nvls_probe.py stand in for them (DESIGN.md section 6)" % (nvlink_measured or {}).get("unavailable")),
                         "nvlink_measured": nvlink_measured,
                         "kernel": ("mt_stream_kernel<DistAdamOp<pair_bf16,pair_bf16,true>>" if p2p else
                                    "mt_stream_kernel<DistAdamOp<mc_bf16,mc_bf16,true>>") + " (includes both barriers)",
                         "bytes_per_launch": nv_bytes, "avg_launch_ms": ms,
                         "peak_source": "measured peer-copy bandwidth per direction (B200_PROFILING.md), 900 nominal",
                         "hbm_gbs_of_shard_kernel": achieved, "hbm_frac": achieved / peak},
            "cpu_baseline": None,
            "e2e": {"value": nparams / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 2 * total * world, "d2h_bytes_per_step": 2 * world,
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "pinned_on_gpu_numa_node": numa_known, "host_numa_nodes": host_numa_nodes(),
                    "note": "every rank copies its whole 2-byte gradient buffer from host memory each step (data-parallel ranks all hold full "
                            "gradients): N x 16 GB per step over N PCIe links out of ONE host memory system, so this number is paced by "
                            "PCIe and host DRAM, not by the optimizer",
                    "api": "DistributedFusedAdam(fused_collectives=True).complete_reductions() + step(), grads copied from pinned host each step"},
            "gpu_launches": int(tl.item()),
            "clocks": clocks,
            "parity": parity,
        }
        print(json.dumps(line))
    dist.destroy_process_group()
    if not parity_ok:
        raise SystemExit("bench_sharded: multi-GPU parity bound violated (see the `parity` block of the JSON line)")


def run_sharded(args):
    from bench import ClockSampler, peaks, METRIC, UNIT, ROOT
    from apex.contrib.optimizers import DistributedFusedAdam
    import _b200mt as mt

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    model_name = os.environ.get("B200MT_SHARDED_MODEL", "llama3_8b")
    K = int(os.environ.get("B200MT_SHARDED_CHUNKS", "8"))
    # default: the whole step as one multimem kernel (DistributedFusedAdam(fused_collectives=True));
    # B200MT_SHARDED_FUSED=0 times the NCCL pipeline RS -> Adam -> AG instead (profiles/sharded_bench_r01.jsonl has both)
    fused = os.environ.get("B200MT_SHARDED_FUSED", "1") == "1"
    if fused:
        K = 1
    shapes = param_shapes.MODELS[model_name]()
    nparams = param_shapes.numel(shapes)

    # B200MT_SHARDED_E5M2=1 (not the headline): the reference's e5m2_allgather=True -- parameters travel as one byte
    # each and are unpacked into the bf16 model parameters by one extra maybe_cast_mt launch (no zero-copy then)
    e5m2 = os.environ.get("B200MT_SHARDED_E5M2", "0") == "1"

    def make(fused_flag, chunks):
        """Model parameters (bf16, seed 1 on every rank), the optimizer, and its flat gradient buffer filled with
        randn(seed = 100 + rank): the same state whichever step implementation is built on it."""
        gen = torch.Generator(device=dev).manual_seed(1)
        params = []
        for s in shapes:
            params.append(torch.nn.Parameter((torch.randn(s, generator=gen, device=dev, dtype=torch.float32) * 0.02).to(torch.bfloat16)))
        o = DistributedFusedAdam(params, lr=1e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, dwu_num_blocks=chunks,
                                 dwu_num_chunks=1, predivide=False, overlap_reductions=False, zero_copy_params=not e5m2,
                                 fused_collectives=((os.environ.get("B200MT_SHARDED_FORM") or True) if fused_flag else False),
                                 e5m2_allgather=e5m2)
        o.set_global_scale(float(world))        # gradients are summed by the reduce-scatter; the kernel divides by W
        fl = o.attach_flat_grads()
        ggen = torch.Generator(device=dev).manual_seed(100 + rank)
        piece = 1 << 26
        for off in range(0, fl.numel(), piece):   # randn(seed = rank) in pieces (fp32 staging)
            n = min(piece, fl.numel() - off)
            fl[off:off + n].copy_(torch.randn(n, generator=ggen, device=dev))
        torch.cuda.synchronize()
        return o, fl

    def make_fused_or_none():
        try:
            o, fl = make(True, 1)
            ok = torch.ones(1, device=dev)
        except Exception as e:                      # no multicast-capable symmetric memory on this box
            sys.stderr.write("bench_sharded: fused multimem path unavailable (%r)\n" % (e,))
            ok, o, fl = torch.zeros(1, device=dev), None, None
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)   # every rank switches together
        return (o, fl) if float(ok) != 0.0 else (None, None)

    # ---- parity before anything is timed (tools/multi_gpu_parity.py): one step of BOTH step implementations from
    # identical state, each compared with the unsharded single-GPU kernel on a sample of every rank's shard and on the
    # gathered parameters of every rank; then the sharded LAMB and the DDP bucket path on small models.  A violated bound
    # makes the run exit non-zero after the JSON line is printed.
    parity, parity_ok = None, True
    if os.environ.get("B200MT_SHARDED_PARITY", "1") == "1":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import multi_gpu_parity as mgp
        parity = {"reference_test": "tests/L0/run_optimizers/test_dist_adam.py:119-173 (N-rank DistributedFusedAdam vs the unsharded optimizer)",
                  "unsharded_kernel": "b200mt_dist_adam on the same reduced gradient (pinned to the reference's kernel by tests/test_parity_gpu.py)"}
        o1, _ = make(False, K if not fused else int(os.environ.get("B200MT_SHARDED_CHUNKS", "8")))
        parity["nccl_vs_unsharded"], ok1 = mgp.sharded_adam_parity(o1)
        del o1, _
        import gc
        gc.collect(); torch.cuda.empty_cache()
        ok2 = True
        if fused:
            opt, flat = make_fused_or_none()
            if opt is not None:
                parity["fused_vs_unsharded"], ok2 = mgp.sharded_adam_parity(opt)
            else:
                fused = False
                parity["fused_vs_unsharded"] = "multicast memory unavailable"
        parity["sharded_lamb_vs_unsharded_f64"], ok3 = mgp.sharded_lamb_parity(dev, rank, world)
        parity["ddp_bucket_path_vs_mean_gradient"], ok4 = mgp.ddp_parity(dev, rank, world)
        parity_ok = bool(ok1 and ok2 and ok3 and ok4)
        parity["ok"] = parity_ok
        if not fused:
            K = int(os.environ.get("B200MT_SHARDED_CHUNKS", "8"))
            opt, flat = make(False, K)
    else:
        opt = None
        if fused:
            opt, flat = make_fused_or_none()
            if opt is None:
                fused = False
        if opt is None:
            K = int(os.environ.get("B200MT_SHARDED_CHUNKS", "8"))
            opt, flat = make(False, K)
    torch.cuda.synchronize()
    lib = mt.lib

    def step():
        opt.complete_reductions()
        opt.step()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    stream = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = lib.b200mt_launch_count()
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from nvlink_counters import NvlinkCounters
    nvl = NvlinkCounters(local)
    torch.cuda.synchronize()
    dist.barrier()
    nv0 = nvl.read() if nvl.ok else None
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    dist.barrier()
    ms = e0.elapsed_time(e1) / args.steps
    # bytes this GPU's NVLinks carried per step, from the driver's counters (max over ranks below)
    nv_tx, nv_rx = ([d / args.steps for d in nvl.delta(nv0)] if nvl.ok else (None, None))
    tnv = torch.tensor([nv_tx or 0.0, nv_rx or 0.0, 1.0 if nvl.ok else 0.0], device=dev, dtype=torch.float64)
    dist.all_reduce(tnv, op=dist.ReduceOp.MAX)
    nvlink_measured = None
    if float(tnv[2]) > 0:
        nvlink_measured = {"tx_bytes_per_step_per_gpu": float(tnv[0]), "rx_bytes_per_step_per_gpu": float(tnv[1]),
                           "tx_gbs": float(tnv[0]) / (ms * 1e-3) / 1e9, "rx_gbs": float(tnv[1]) / (ms * 1e-3) / 1e9,
                           "source": "NVML NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX/RX summed over the GPU's links, read around the "
                                     "timed loop; max over ranks"}
    else:
        nvlink_measured = {"unavailable": nvl.err}
    launches = lib.b200mt_launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None

    shard_elems = opt._layout["shard"] * K
    if fused:
        return _report_fused(args, opt, ms, launches, sampler, rank, world, local, dev, stream, flat, nparams, shapes, model_name, lib,
                             parity, parity_ok, nvlink_measured)
    # shard kernel alone (no collectives): the roofline of the dominant kernel on this rank
    ks = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    kreps = 5
    torch.cuda.synchronize()
    ks[0].record(stream)
    for _ in range(kreps):
        for k in range(K):
            opt._adam_chunk(k, float(world))
    ks[1].record(stream)
    torch.cuda.synchronize()
    kernel_ms = ks[0].elapsed_time(ks[1]) / kreps

    # collectives alone
    cs = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    chunk, shard = opt._layout["chunk"], opt._layout["shard"]
    torch.cuda.synchronize(); dist.barrier()
    cs[0].record(stream)
    for k in range(K):
        dist.reduce_scatter_tensor(opt._fp16_g[k * shard:(k + 1) * shard], opt._flat_grads[k * chunk:(k + 1) * chunk])
    cs[1].record(stream)
    for k in range(K):
        dist.all_gather_into_tensor(opt._new_params[k * chunk:(k + 1) * chunk], opt._fp16_p[k * shard:(k + 1) * shard])
    cs[2].record(stream)
    torch.cuda.synchronize()
    rs_ms, ag_ms = cs[0].elapsed_time(cs[1]), cs[1].elapsed_time(cs[2])

    # e2e: the step's gradients arrive from pinned host memory (a 256 MiB pinned window copied repeatedly to
    # fill the flat buffer: pinning 16 GB per rank is not reasonable), the result read back is the gradient norm
    from bench import pinned_near_gpu, gpu_numa_cpus, host_numa_nodes
    host = pinned_near_gpu(1 << 27, torch.bfloat16, local)
    numa_known = gpu_numa_cpus(local) is not None or host_numa_nodes() == 1
    host.normal_(0, 1)
    e2e_steps = max(2, min(args.steps, 5))

    def e2e_step():
        for off in range(0, flat.numel(), host.numel()):
            n = min(host.numel(), flat.numel() - off)
            flat[off:off + n].copy_(host[:n], non_blocking=True)
        opt.complete_reductions()
        opt.step()
        return float(opt._fp16_p[0])     # D2H read of one updated parameter: forces completion of the step

    e2e_step()
    torch.cuda.synchronize(); dist.barrier()
    e0.record(stream)
    for _ in range(e2e_steps):
        e2e_step()
    e1.record(stream)
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps

    t = torch.tensor([ms, kernel_ms, rs_ms, ag_ms, e2e_ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tl = torch.tensor([float(launches)], device=dev, dtype=torch.float64)
    dist.all_reduce(tl, op=dist.ReduceOp.SUM)
    ms, kernel_ms, rs_ms, ag_ms, e2e_ms = t.tolist()
    if rank == 0:
        peak, peak_src = peaks()
        total = opt._layout["total"]
        pbytes = 1.0 if opt._e5m2 else 2.0
        kbytes = (26.0 + pbytes) * shard_elems
        achieved = kbytes / (kernel_ms * 1e-3) / 1e9
        nv_bytes = 2.0 * total * (world - 1) / world            # per direction per GPU, for the RS (the AG moves pbytes/2 of it)
        ag_bytes = pbytes * total * (world - 1) / world
        line = {
            "metric": METRIC, "value": nparams / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[3]: DistributedFusedAdam sharded step on the %s shape list: bf16 grad reduce-scatter -> "
                                   "shard-local fused AdamW (fp32 master/moments, bf16 copy) -> bf16 param all-gather, NCCL" % model_name,
                       "tensors": len(shapes), "params": nparams, "flat_elems": total, "pipeline_chunks": K,
                       "parallelism": "zero-dp%d" % world,
                       "l2": "inputs larger than L2 (%.1f GB of optimizer state per rank)" % (12.0 * shard_elems / 1e9)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "kernel": "mt_stream_kernel<DistAdamOp<__nv_bfloat16,true>>", "bytes_per_launch": kbytes / K,
                         "avg_launch_ms": kernel_ms / K, "peak_source": peak_src},
            "nvlink": {"bytes_per_direction_per_gpu_per_collective": nv_bytes, "reduce_scatter_ms": rs_ms, "all_gather_ms": ag_ms,
                       "reduce_scatter_gbs": nv_bytes / (rs_ms * 1e-3) / 1e9, "all_gather_gbs": ag_bytes / (ag_ms * 1e-3) / 1e9,
                       "peak_gbs": NVLINK_GBS, "frac_rs": nv_bytes / (rs_ms * 1e-3) / 1e9 / NVLINK_GBS,
                       "frac_ag": ag_bytes / (ag_ms * 1e-3) / 1e9 / NVLINK_GBS,
                       "step_floor_ms": (nv_bytes + ag_bytes) / (NVLINK_GBS * 1e9) * 1e3, "e5m2_allgather": bool(opt._e5m2),
                       "overlap_efficiency": (rs_ms + ag_ms) / ms, "shard_kernel_ms": kernel_ms, "measured": nvlink_measured},
            "cpu_baseline": None,
            "e2e": {"value": nparams / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 2 * total * world, "d2h_bytes_per_step": 2 * world,
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "pinned_on_gpu_numa_node": numa_known,
                    "api": "DistributedFusedAdam.complete_reductions() + step(), grads copied from pinned host each step"},
            "gpu_launches": int(tl.item()),
            "clocks": clocks,
            "parity": parity,
        }
        print(json.dumps(line))
    dist.destroy_process_group()
    if not parity_ok:
        raise SystemExit("bench_sharded: multi-GPU parity bound violated (see the `parity` block of the JSON line)")
