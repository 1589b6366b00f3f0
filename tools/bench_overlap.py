"""Measurement (not product): which sharded step should a trainer use when there is a backward pass to hide behind?

  torchrun --nproc-per-node N tools/bench_overlap.py [--backward-ms 40]

One iteration = a synthetic backward of known length (a chain of bf16 matmuls on the main stream, cut into as many slices as
the optimizer has pipeline chunks; after each slice the gradients of one chunk are declared produced, last chunk first, the order
backward produces them) followed by complete_reductions() + step().  Llama-3-8B shape list, bf16 gradients, fp32 masters.

  fused     DistributedFusedAdam(fused_collectives=True): nothing can start before the last gradient exists; the whole step
            is one multimem kernel after backward.
  pipeline  DistributedFusedAdam(overlap_reductions=True, K chunks): the gradient hooks start a chunk's reduce-scatter on the
            side stream while backward goes on; after backward only the last reduce-scatter, the shard Adam and the
            all-gathers remain (the reference's design, apex/contrib/optimizers/distributed_fused_adam.py:397-477,489-502).

Reported per mode: iteration time, the backward alone, and their difference = what the optimizer adds to the iteration.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist

import param_shapes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backward-ms", type=float, default=40.0)
    ap.add_argument("--chunks", type=int, default=8)
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from apex.contrib.optimizers import DistributedFusedAdam
    shapes = param_shapes.llama3_8b()
    nparams = param_shapes.numel(shapes)

    # ---- the synthetic backward: n matmuls of 4096^3 bf16, calibrated to --backward-ms
    a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    b = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    c = torch.empty_like(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(10):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(200):
        torch.matmul(a, b, out=c)
    e1.record()
    torch.cuda.synchronize()
    per_mm = e0.elapsed_time(e1) / 200

    def make(fused):
        gen = torch.Generator(device=dev).manual_seed(1)
        params = [torch.nn.Parameter((torch.randn(s, generator=gen, device=dev) * 0.02).to(torch.bfloat16)) for s in shapes]
        K = 1 if fused else args.chunks
        opt = DistributedFusedAdam(params, lr=1e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, dwu_num_blocks=K, dwu_num_chunks=1,
                                   predivide=False, overlap_reductions=not fused, zero_copy_params=True, fused_collectives=fused)
        opt.set_global_scale(float(world))
        flat = opt.attach_flat_grads()
        flat.normal_(0, 1)
        return opt

    def run(opt, fused):
        K = opt._num_chunks_total
        slices = args.chunks
        mm_per_slice = max(1, int(round(args.backward_ms / slices / per_mm)))

        def backward(with_hooks):
            for s in range(slices):
                for _ in range(mm_per_slice):
                    torch.matmul(a, b, out=c)
                if with_hooks and not fused:
                    for i in opt._chunk_params[K - 1 - s]:   # chunks become complete last to first (slices == K)
                        opt._on_grad(i)

        def iteration():
            backward(True)
            opt.complete_reductions()
            opt.step()

        for _ in range(2):
            iteration()
        torch.cuda.synchronize(); dist.barrier()
        e0.record()
        for _ in range(args.iters):
            backward(False)
        e1.record()
        torch.cuda.synchronize()
        bwd = e0.elapsed_time(e1) / args.iters
        dist.barrier()
        e0.record()
        for _ in range(args.iters):
            iteration()
        e1.record()
        torch.cuda.synchronize()
        it = e0.elapsed_time(e1) / args.iters
        # one instrumented iteration: when, relative to the end of backward, are the reduce-scatters done and the step over?
        phases = None
        if not fused:
            eb, er, es = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            torch.cuda.synchronize(); dist.barrier()
            e0.record()
            backward(True)
            eb.record()                                   # end of backward on the main stream
            opt.complete_reductions()
            er.record(opt._comm_stream)                   # every reduce-scatter issued so far has completed
            opt.step()
            es.record()
            torch.cuda.synchronize()
            phases = {"backward_end_ms": round(e0.elapsed_time(eb), 2), "reduce_scatters_done_ms": round(e0.elapsed_time(er), 2),
                      "step_done_ms": round(e0.elapsed_time(es), 2)}
        t = torch.tensor([bwd, it], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), phases

    out = {"world": world, "params": nparams, "backward_ms_requested": args.backward_ms, "matmul_4096_bf16_ms": per_mm}
    for name, fused in (("pipeline_overlapped", False), ("fused_after_backward", True)):
        try:
            opt = make(fused)
            bwd, it, phases = run(opt, fused)
            out[name] = {"backward_ms": round(bwd, 2), "iteration_ms": round(it, 2), "optimizer_adds_ms": round(it - bwd, 2)}
            if phases:
                out[name]["one_iteration_timeline"] = phases
            del opt
            import gc
            gc.collect(); torch.cuda.empty_cache()
        except Exception as e:       # e.g. no multicast memory on this box
            out[name] = {"error": repr(e)}
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
