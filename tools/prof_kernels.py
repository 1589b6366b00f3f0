"""Profiling target: a few launches of the non-headline kernels on the BERT-large list (l2norm_mp, lamb_mp stage 1/2 +
norm second pass, mixed-precision Adam, sharded-Adam kernel), for `ncu --set full`."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import amp_C, distributed_adam_cuda, param_shapes
dev = torch.device("cuda:0")
shapes = param_shapes.bert_large()
g = torch.Generator(device=dev).manual_seed(0)
mk = lambda dt, s: [(torch.randn(sh, generator=g, device=dev) * s).to(dt) for sh in shapes]
g16, p, m, v = mk(torch.float16, 1.0), mk(torch.float32, 0.02), mk(torch.float32, 0.01), [t.abs() for t in mk(torch.float32, 0.001)]
p16 = [t.half() for t in p]
gb, pb = [t.bfloat16() for t in g16], [t.bfloat16() for t in p]
f = torch.zeros(1, dtype=torch.int32, device=dev)
one = lambda val, dt=torch.float32: torch.tensor([val], dtype=dt, device=dev)
lr, step, inv_scale, found_inf, max_norm = one(1e-3), one(5, torch.int32), one(1.0), one(0.0), one(1e9)
T = len(shapes)
b1, b2, eps, wd = torch.full((T,), 0.9, device=dev), torch.full((T,), 0.999, device=dev), torch.full((T,), 1e-8, device=dev), torch.full((T,), 0.01, device=dev)
bc = torch.ones(T, dtype=torch.int32, device=dev)
for it in range(3):
    gn, _ = amp_C.multi_tensor_l2norm_mp(65536, f, [g16], False)
    amp_C.multi_tensor_lamb_mp(65536, f, [g16, p, m, v, p16], lr, 0.9, 0.999, 1e-6, step, 1, 0.01, 1, 1, gn, max_norm, False, found_inf, inv_scale)
    amp_C.multi_tensor_adam(65536, f, [g16, p, m, v, p16], 1e-3, 0.9, 0.999, 1e-8, it + 1, 1, 1, 0.0)
    distributed_adam_cuda.multi_tensor_fused_adam(65536, f, [p, m, v, gb, pb], b1, b2, bc, eps, wd, 1e-3, 1.0, it + 1, 1)
torch.cuda.synchronize()
print("done")
