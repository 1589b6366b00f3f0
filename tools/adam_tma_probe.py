"""Probe: fp32 multi_tensor_adam with direct 128-bit loads/stores (default) or with TMA bulk-copy staging
(B200MT_ADAM_TMA=1, csrc/ops_adam_tma.cu).  Checks the result against a float64 torch evaluation of the rule and times
back-to-back launches on the ResNet-50 and BERT-large lists.  One B200; run once per setting."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import amp_C, param_shapes
import _b200mt as mt

dev = torch.device("cuda:0")
f = torch.zeros(1, dtype=torch.int32, device=dev)
out = {"tma": os.environ.get("B200MT_ADAM_TMA", "0"), "tile": os.environ.get("B200MT_TILE_ELEMS", "4096")}

# correctness on ragged sizes (tails of 1..3 elements, tiny tensors, several tiles)
g = torch.Generator(device=dev).manual_seed(0)
sizes = [1, 2, 3, 5, 4096, 4097, 8191, 12289, 65537, 100003]
mk = lambda s: [torch.randn(n, generator=g, device=dev) * s for n in sizes]
gr, p, m, v = mk(1.0), mk(0.1), mk(0.01), [t.abs() for t in mk(0.001)]
ref = [[t.double().clone() for t in l] for l in (gr, p, m, v)]
lr, b1, b2, eps, step, wd = 1e-2, 0.9, 0.99, 1e-8, 3, 0.01
for mode in (1, 0):
    amp_C.multi_tensor_adam(65536, f, [gr, p, m, v], lr, b1, b2, eps, step, mode, 1, wd)
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    for G, P, M, V in zip(*ref):
        gg = G + wd * P if mode == 0 else G
        M.mul_(b1).add_(gg, alpha=1 - b1)
        V.mul_(b2).add_(gg * gg, alpha=1 - b2)
        upd = (M / bc1) / ((V / bc2).sqrt() + eps)
        P.sub_(lr * (upd + (wd * P if mode == 1 else 0)))
    err = max(float((a.double() - b).abs().max()) for l, r in zip((p, m, v), ref[1:]) for a, b in zip(l, r))
    out["max_abs_err_mode%d" % mode] = err
    assert err < 1e-6, err

for name in ("resnet50", "bert_large"):
    shapes = param_shapes.MODELS[name]()
    mk = lambda s: [torch.randn(sh, generator=g, device=dev) * s for sh in shapes]
    gr, p, m, v = mk(1e-3), mk(0.02), mk(0.0), [t.abs() for t in mk(0.0)]
    n = sum(t.numel() for t in p)
    for _ in range(5):
        amp_C.multi_tensor_adam(65536, f, [gr, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 5, 1, 1, 0.0)
    torch.cuda.synchronize()
    reps = 100 if name == "resnet50" else 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = mt.lib.b200mt_launch_count()
    e0.record()
    for _ in range(reps):
        amp_C.multi_tensor_adam(65536, f, [gr, p, m, v], 1e-3, 0.9, 0.999, 1e-8, 5, 1, 1, 0.0)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    out[name] = {"us": round(us, 1), "GBs": round(28.0 * n / us / 1e3), "launches": int(mt.lib.b200mt_launch_count() - l0)}
print(json.dumps(out))
