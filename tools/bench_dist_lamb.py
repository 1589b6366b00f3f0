"""Measurement tool (not the driver's bench): the sharded LAMB step of DistributedFusedLAMB on the BERT-large
(default) or Llama-3-8B shape list, one process per GPU under torch.distributed.run."""
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
    from apex.contrib.optimizers import DistributedFusedLAMB
    import _b200mt as mt
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    model = os.environ.get("MODEL", "bert_large")
    shapes = param_shapes.MODELS[model]()
    n = param_shapes.numel(shapes)
    gen = torch.Generator(device=dev).manual_seed(1)
    params = [torch.nn.Parameter((torch.randn(s, generator=gen, device=dev) * 0.02).to(torch.bfloat16)) for s in shapes]
    opt = DistributedFusedLAMB(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, max_grad_norm=1.0,
                               overlap_reductions=False, zero_copy_params=True)
    opt.set_global_scale(1.0)
    flat = opt.attach_flat_grads()
    flat.normal_(0, 1)
    torch.cuda.synchronize()

    def step():
        opt.complete_reductions()
        opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = mt.lib.b200mt_launch_count()
    steps = 10
    e0.record()
    for _ in range(steps):
        step()
    e1.record(); torch.cuda.synchronize()
    launches = (mt.lib.b200mt_launch_count() - l0) / steps
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # the two kernels alone on this rank's shard
    ks = []
    h = opt._hyper
    for fn in (lambda: opt.multi_tensor_lamb_compute_update_term(65536, opt._overflow_buf, opt._stage1_lists, h["beta1"], h["beta2"], h["beta3"],
                                                                 h["bias_correction"], opt._step_t, h["eps"], 1, h["decay"], opt._global_scale_t,
                                                                 opt._L2_grad_norm, 1.0, opt._frag_sumsq),
               lambda: opt.multi_tensor_lamb_update_weights(65536, opt._overflow_buf, opt._stage2_lists, torch.ones(len(opt._stage2_lists[0]), device=dev),
                                                            torch.ones(len(params), device=dev), opt._frag_param, opt._lr, h["decay"], opt._L2_grad_norm, False)):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            fn()
        b.record(); torch.cuda.synchronize()
        ks.append(a.elapsed_time(b) / 5)
    if rank == 0:
        shard = opt._layout["shard"]
        print(json.dumps({"op": "DistributedFusedLAMB step", "model": model, "n_gpus": world, "params": n, "ms_per_step": float(t),
                          "params_per_s": n / float(t) * 1e3, "our_launches_per_step": launches,
                          "stage1_ms": ks[0], "stage1_GBs(2+12r,12w=26B)": 26.0 * shard / ks[0] / 1e6,
                          "stage2_ms": ks[1], "stage2_GBs(8r,6w=14B)": 14.0 * shard / ks[1] / 1e6}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
