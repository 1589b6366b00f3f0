"""Probe: tile size for the norm-type kernels (l2norm, lamb stage 1 with fused norms), one B200."""
import os, sys, statistics, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import _b200mt as mt, amp_C, param_shapes
dev = torch.device("cuda:0")
flush = torch.empty(64 << 20, device=dev)
def timeit(fn, reps=7):
    fn(); fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)
f = torch.zeros(1, dtype=torch.int32, device=dev)
big = torch.randn(4 * (1 << 28), device=dev)
xs = list(big.view(4, -1).unbind(0))
half = big[: 1 << 29].half(); xh = list(half.view(4, -1).unbind(0))
shapes = param_shapes.bert_large()
g = torch.Generator(device=dev).manual_seed(0)
mk = lambda dt, s: [(torch.randn(sh, generator=g, device=dev) * s).to(dt) for sh in shapes]
g16, p, m, v = mk(torch.float16, 1.0), mk(torch.float32, 0.02), mk(torch.float32, 0.01), [t.abs() for t in mk(torch.float32, 0.001)]
p16 = [t.half() for t in p]
one = lambda val, dt=torch.float32: torch.tensor([val], dtype=dt, device=dev)
lr, step, inv_scale, found_inf, max_norm, gn = one(1e-3), one(5, torch.int32), one(1.0), one(0.0), one(1e9), one(1.0)
for tile in (2048, 4096, 8192, 16384, 32768, 65536):
    mt.check(mt.lib.b200mt_set_tuning(tile, 0))
    a = timeit(lambda: amp_C.multi_tensor_l2norm(65536, f, [xs], True))
    b = timeit(lambda: amp_C.multi_tensor_l2norm(65536, f, [xh], False))
    c = timeit(lambda: amp_C.multi_tensor_lamb_mp(65536, f, [g16, p, m, v, p16], lr, 0.9, 0.999, 1e-6, step, 1, 0.01, 1, 1, gn, max_norm, False, found_inf, inv_scale))
    d = timeit(lambda: amp_C.multi_tensor_scale(65536, f, [xh, xs[:4]], 0.5) if False else amp_C.multi_tensor_scale(65536, f, [xs, xs], 1.0))
    print(json.dumps({"tile": tile, "l2norm_fp32_4x256M_us": round(a, 1), "GBs": round(4 * (1 << 30) / a / 1e3), "l2norm_fp16_4x128M_us": round(b, 1),
                      "GBs16": round((1 << 30) / b / 1e3), "lamb_mp_bert_us": round(c, 1), "lamb_GBs(38B)": round(38 * 336226108 / c / 1e3),
                      "scale_fp32_inplace_us": round(d, 1)}), flush=True)
