"""Probe (not product): where does the fused multimem step spend its time?  One torchrun process per GPU;
times, on 16 GB flat bf16 buffers in symmetric memory, (a) the stand-alone multimem reduce-scatter kernel,
(b) the stand-alone multimem all-gather kernel, (c) the fused Adam step, each for a few tile sizes / grid caps."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "rocm-apex_b200"))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

import _b200mt as mt


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    total = int(os.environ.get("PROBE_ELEMS", str(8030261248)))
    total = total // (256 * world) * (256 * world)
    shard = total // world
    bf = torch.bfloat16
    g = symm_mem.empty(total, dtype=bf, device=dev)
    p = symm_mem.empty(total, dtype=bf, device=dev)
    g.normal_(0, 1); p.zero_()
    hg = symm_mem.rendezvous(g, dist.group.WORLD.group_name)
    hp = symm_mem.rendezvous(p, dist.group.WORLD.group_name)
    out = torch.zeros(shard, dtype=bf, device=dev)
    p32 = torch.zeros(shard, device=dev); m32 = torch.zeros(shard, device=dev); v32 = torch.zeros(shard, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    off = rank * shard * 2
    g_mc = mt.RawBuffer(hg.multicast_ptr + off, shard, bf, dev)
    p_mc = mt.RawBuffer(hp.multicast_ptr + off, shard, bf, dev)
    L_rs = mt.pack([[g_mc], [out]], 2)
    L_ag = mt.pack([[out], [p_mc]], 2)
    L_fu = mt.pack([[p32], [m32], [v32], [g_mc], [p_mc]], 5)
    one = lambda v, dt=torch.float32: torch.tensor([v], dtype=dt, device=dev)
    b1, b2, ep, wd, bc = one(0.9), one(0.95), one(1e-8), one(0.1), one(1, torch.int32)
    sp = torch.cuda.current_stream(dev).cuda_stream
    lib = mt.lib

    def rs():
        hg.barrier(channel=0)
        mt.check(lib.b200mt_nvls_reduce_scatter(L_rs.ref(), 65536, sp))
        hg.barrier(channel=1)

    def ag():
        hp.barrier(channel=0)
        mt.check(lib.b200mt_nvls_all_gather(L_ag.ref(), 65536, sp))
        hp.barrier(channel=1)

    step = [0]

    def fused():
        step[0] += 1
        hg.barrier(channel=0)
        mt.check(lib.b200mt_dist_adam_nvls(L_fu.ref(), 65536, flag.data_ptr(), b1.data_ptr(), b2.data_ptr(), bc.data_ptr(),
                                           ep.data_ptr(), wd.data_ptr(), 1e-4, float(world), step[0], 1, sp))
        hp.barrier(channel=1)

    # the fused kernel with only ONE of its two collectives (include/b200mt.h: B200MT_NVLS_LOCAL_COPY / _LOCAL_GRADS)
    g_loc = g[rank * shard:(rank + 1) * shard]
    L_fa = mt.pack([[p32], [m32], [v32], [g_mc], [out]], 5)       # reduce in the switch, store locally
    L_fb = mt.pack([[p32], [m32], [v32], [g_loc], [p_mc]], 5)     # load locally, broadcast through the switch

    def fused_form(L, flag_bits):
        def run():
            step[0] += 1
            hg.barrier(channel=0)
            mt.check(lib.b200mt_dist_adam_nvls(L.ref(), 65536, flag.data_ptr(), b1.data_ptr(), b2.data_ptr(), bc.data_ptr(),
                                               ep.data_ptr(), wd.data_ptr(), 1e-4, float(world), step[0], 1 | flag_bits, sp))
            hp.barrier(channel=1)
        return run

    fused_reduce_only, fused_bcast_only = fused_form(L_fa, 0x200), fused_form(L_fb, 0x100)
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from nvlink_counters import NvlinkCounters
    nvl = NvlinkCounters(dev.index)

    def timed(fn, reps=4):
        fn(); fn(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    # ---- which direction paces the fused step?  three single launches + the links' own byte counters
    forms = {}
    for name, fn in (("full", fused), ("reduce_only_local_store", fused_reduce_only), ("broadcast_only_local_load", fused_bcast_only),
                     ("standalone_rs", rs), ("standalone_ag", ag)):
        reps = 4
        fn(); fn(); torch.cuda.synchronize(); dist.barrier()
        before = nvl.read() if nvl.ok else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize(); dist.barrier()
        d = nvl.delta(before) if nvl.ok else [0, 0]
        t = torch.tensor([e0.elapsed_time(e1) / reps, d[0] / reps, d[1] / reps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        forms[name] = {"ms": round(ms, 3), "tx_GB": round(float(t[1]) / 1e9, 3), "rx_GB": round(float(t[2]) / 1e9, 3),
                       "tx_GBs": round(float(t[1]) / ms / 1e6, 1), "rx_GBs": round(float(t[2]) / ms / 1e6, 1)}
    if rank == 0:
        print(json.dumps({"world": world, "elems": total, "model_out_GB_full": round((2.0 * total + 2.0 * shard) / 1e9, 3),
                          "nvlink_counters": nvl.ok, "forms": forms}), flush=True)
    if os.environ.get("PROBE_FORMS_ONLY"):
        dist.destroy_process_group()
        return
    res = []
    for tile, cap in [(4096, 0), (16384, 0), (65536, 0), (262144, 0), (1048576, 0), (16384, 148 * 2), (65536, 148 * 2)]:
        mt.check(lib.b200mt_set_tuning(tile, cap))
        r = {"tile": tile, "grid_cap": cap, "rs_ms": timed(rs), "ag_ms": timed(ag), "fused_ms": timed(fused)}
        r["rs_out_gbs"] = 2.0 * total / r["rs_ms"] / 1e6       # bytes the switch pulls out of each GPU
        r["ag_in_gbs"] = 2.0 * total / r["ag_ms"] / 1e6        # bytes the switch writes into each GPU
        res.append(r)
        if rank == 0:
            print(json.dumps(r), flush=True)
    # correctness spot check of the stand-alone kernels
    mt.check(lib.b200mt_set_tuning(4096, 0))
    rs(); torch.cuda.synchronize()
    want = torch.zeros(4096, device=dev)
    lst = [torch.empty(4096, dtype=bf, device=dev) for _ in range(world)]
    dist.all_gather(lst, g[0:4096].clone())      # everybody's copy of the head of rank 0's shard
    for t in lst:
        want += t.float()
    err = float((out[:4096].float() - want).abs().max() / want.abs().max()) if rank == 0 else 0.0
    if rank == 0:
        print(json.dumps({"world": world, "elems": total, "rs_check_rel_err": err}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
