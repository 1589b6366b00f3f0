"""Probe: the Adam-family kernels of one library build (B200MT_LIB) on the BERT-large list, one B200.
Used to compare the approximate-division build with an IEEE build (-DB200MT_APPROX_MATH=0)."""
import os, sys, statistics, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import amp_C, distributed_adam_cuda, fused_adam_cuda, param_shapes
dev = torch.device("cuda:0")
flush = torch.empty(64 << 20, device=dev)


def timeit(fn, reps=9):
    fn(); fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


shapes = param_shapes.bert_large()
N = sum(int(torch.Size(s).numel()) for s in shapes)
g = torch.Generator(device=dev).manual_seed(0)
mk = lambda dt, s: [(torch.randn(sh, generator=g, device=dev) * s).to(dt) for sh in shapes]
g16, p, m, v = mk(torch.float16, 1.0), mk(torch.float32, 0.02), mk(torch.float32, 0.01), [t.abs() for t in mk(torch.float32, 0.001)]
g32 = [t.float() for t in g16]
p16 = [t.half() for t in p]
gb, pb = [t.bfloat16() for t in g16], [t.bfloat16() for t in p]
f = torch.zeros(1, dtype=torch.int32, device=dev)
one = lambda val, dt=torch.float32: torch.tensor([val], dtype=dt, device=dev)
lr, step, inv_scale, found_inf, max_norm, gn = one(1e-3), one(5, torch.int32), one(1.0), one(0.0), one(1e9), one(1.0)
T = len(shapes)
b1, b2, eps, wd = (torch.full((T,), x, device=dev) for x in (0.9, 0.999, 1e-8, 0.01))
bc = torch.ones(T, dtype=torch.int32, device=dev)
r = {"lib": os.path.basename(os.environ.get("B200MT_LIB", "libb200mt.so")), "params": N}


def rec(name, bytes_per_param, fn):
    us = timeit(fn)
    r[name] = {"us": round(us, 1), "GBs": round(bytes_per_param * N / us / 1e3, 0)}


rec("adam_fp32", 28, lambda: amp_C.multi_tensor_adam(65536, f, [g32, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 5, 1, 1, 0.01))
rec("adam_mixed_g16_p16copy", 28, lambda: amp_C.multi_tensor_adam(65536, f, [g16, p, m, v, p16], 1e-3, 0.9, 0.999, 1e-8, 5, 1, 1, 0.01))
rec("dist_adam_bf16", 28, lambda: distributed_adam_cuda.multi_tensor_fused_adam(65536, f, [p, m, v, gb, pb], b1, b2, bc, eps, wd, 1e-3, 1.0, 5, 1))
rec("contrib_adam_fp32_g16", 26, lambda: fused_adam_cuda.adam_mt(65536, f, [p, m, v, g16], 1e-3, 0.9, 0.999, 1e-8, 1.0, 5, 1, 1, 0.01))
rec("lamb_fp32_3launch", 0, lambda: amp_C.multi_tensor_lamb(65536, f, [g32, p, m, v], 1e-3, 0.9, 0.999, 1e-6, 5, 1, 0.01, 1, 1, gn, 1e9, False))
rec("lamb_mp_3launch", 0, lambda: amp_C.multi_tensor_lamb_mp(65536, f, [g16, p, m, v, p16], lr, 0.9, 0.999, 1e-6, step, 1, 0.01, 1, 1, gn, max_norm, False, found_inf, inv_scale))
rec("adagrad_fp32", 20, lambda: amp_C.multi_tensor_adagrad(65536, f, [g32, p, v], 1e-3, 1e-10, 0, 0.01))
print(json.dumps(r), flush=True)
