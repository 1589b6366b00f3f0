"""Measurement tool: the DDP gradient-bucket path on the ResNet-50 (default) or BERT-large gradient list.
ours      = apex.parallel.DistributedDataParallel's bucket path (fused gather+predivide, all_reduce, fused average+scatter)
reference = the data path of the reference's allreduce_bucket (apex/parallel/distributed.py:426-476) restated with torch ops:
            flatten (cat) -> mul_ -> all_reduce -> mul_ -> unflatten copies, same buckets
Both reduce every gradient once, nothing overlapped with backward; CUDA events, max over ranks."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist

import param_shapes


class Params(torch.nn.Module):
    def __init__(self, shapes, dev):
        super().__init__()
        self.ps = torch.nn.ParameterList([torch.nn.Parameter(torch.randn(s, device=dev) * 0.02) for s in shapes])

    def forward(self, x):
        return sum((p * x).sum() for p in self.ps)


def main():
    from apex.parallel import DistributedDataParallel as DDP
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    model_name = os.environ.get("MODEL", "resnet50")
    shapes = param_shapes.MODELS[model_name]()
    n = param_shapes.numel(shapes)
    model = Params(shapes, dev)
    ddp = DDP(model, message_size=10000000, delay_allreduce=True, gradient_predivide_factor=1.0)
    x = torch.ones((), device=dev)
    ddp(x).backward()                    # builds the buckets
    torch.cuda.synchronize()
    buckets = [b.params for b in ddp._buckets]

    def ours():
        ddp._end_of_backward()

    def reference_style():
        for params in buckets:
            grads = [p.grad for p in params]
            flat = torch.cat([g.reshape(-1) for g in grads])
            flat.mul_(1.0)               # predivide (factor 1.0 is still a pass in the reference when != 1; kept out: see below)
            dist.all_reduce(flat)
            flat.mul_(1.0 / world)
            off = 0
            for g in grads:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()

    def allreduce_only():
        for b in ddp._buckets:
            dist.all_reduce(b.flat)

    res = {"op": "DDP gradient reduction", "model": model_name, "n_gpus": world, "params": n, "buckets": len(buckets)}
    for name, fn in (("ours", ours), ("reference_style", reference_style), ("allreduce_only", allreduce_only)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name + "_ms"] = float(t)
    res["speedup_vs_reference_style"] = res["reference_style_ms"] / res["ours_ms"]
    res["busbw_GBs_ours"] = 4.0 * n * 2 * (world - 1) / world / res["ours_ms"] / 1e6
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
