"""Probe (not product): NCCL reduce-scatter / all-gather bandwidth on the flat bf16 layout, alone and
concurrently on two communicators, and whether torch symmetric memory offers a multicast pointer."""
import os, json
import torch, torch.distributed as dist

def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    pg2 = dist.new_group(backend="nccl")
    total = int(os.environ.get("PROBE_ELEMS", str(8 * 1024**3)))   # 8 Gi bf16 = 16 GiB
    K = int(os.environ.get("PROBE_CHUNKS", "8"))
    total = total // (256 * K * world) * (256 * K * world)
    chunk, shard = total // K, total // K // world
    flat = torch.zeros(total, dtype=torch.bfloat16, device=dev)
    full = torch.zeros(total, dtype=torch.bfloat16, device=dev)
    gs = torch.zeros(K * shard, dtype=torch.bfloat16, device=dev)
    ps = torch.zeros(K * shard, dtype=torch.bfloat16, device=dev)
    s1, s2 = torch.cuda.Stream(priority=-1), torch.cuda.Stream(priority=-1)
    def rs(group=None):
        for k in range(K):
            dist.reduce_scatter_tensor(gs[k*shard:(k+1)*shard], flat[k*chunk:(k+1)*chunk], group=group)
    def ag(group=None):
        for k in range(K):
            dist.all_gather_into_tensor(full[k*chunk:(k+1)*chunk], ps[k*shard:(k+1)*shard], group=group)
    def timed(fn, reps=3):
        fn(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)
    def both():
        cur = torch.cuda.current_stream()
        s1.wait_stream(cur); s2.wait_stream(cur)
        with torch.cuda.stream(s1): rs(None)
        with torch.cuda.stream(s2): ag(pg2)
        cur.wait_stream(s1); cur.wait_stream(s2)
    res = {"world": world, "elems": total, "K": K}
    res["rs_ms"] = timed(rs); res["ag_ms"] = timed(ag); res["rs_and_ag_concurrent_ms"] = timed(both)
    nb = 2.0 * total * (world - 1) / world
    res["rs_gbs"] = nb / res["rs_ms"] / 1e6; res["ag_gbs"] = nb / res["ag_ms"] / 1e6
    res["concurrent_gbs_each"] = nb / res["rs_and_ag_concurrent_ms"] / 1e6
    # symmetric memory / multicast
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(1 << 20, dtype=torch.bfloat16, device=dev)
        h = symm.rendezvous(t, dist.group.WORLD.group_name)
        res["symm_backend"] = str(symm.get_backend(dev)) if hasattr(symm, "get_backend") else None
        res["multicast_ptr"] = int(h.multicast_ptr)
        res["n_buffer_ptrs"] = len(h.buffer_ptrs)
        try:
            res["has_multicast_support"] = bool(type(h).has_multicast_support(torch._C._autograd.DeviceType.CUDA if False else h.device.type if False else dist.distributed_c10d.DeviceType.CUDA, dev.index))
        except Exception as e:
            res["has_multicast_support"] = "n/a: " + str(e)[:80]
        h.barrier(0)
        res["symm_ok"] = True
    except Exception as e:
        res["symm_ok"] = False; res["symm_err"] = repr(e)[:300]
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()

if __name__ == "__main__":
    main()
