"""Probe (not product): one library build (B200MT_LIB) on the light streaming ops the round-1 verdict found furthest
below the roofline -- cold-cache (L2 flushed before every timed launch) per-call time through the C ABI, one B200:

  ResNet-50 list   scale fp32 in place, adam fp32, and the warm back-to-back step (scale + adam) of bench.py
  1000 x 64K       scale, l2norm, axpby           100 x 1M   l2norm          BERT-large  l2norm fp16 (the LAMB grad norm)
  10000 x 1K       scale / l2norm / axpby through amp_C.* (host-bound regime: wall time per call, no flush)

Prints one JSON line; profiles/sweep_light_r02.txt collects them per build."""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch

import _b200mt as mt
import amp_C
import param_shapes

dev = torch.device("cuda:0")
lib = mt.lib
CHUNK = 65536
PEAK = 6548.2
flush = torch.empty(128 << 20, device=dev)
stream = torch.cuda.current_stream(dev)
sp = stream.cuda_stream
f = torch.zeros(1, dtype=torch.int32, device=dev)
fp = f.data_ptr()


def cold(call, reps=15):
    call(); call()
    ts = []
    for k in range(reps):
        flush.fill_(float(k))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); mt.check(call()); e1.record(stream)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


def views(T, n, dtype=torch.float32):
    big = torch.randn(T * n, device=dev).to(dtype)
    return big, list(big.view(T, n).unbind(0))


r = {"lib": os.path.basename(os.environ.get("B200MT_LIB", "libb200mt.so"))}

# ---- ResNet-50
shapes = param_shapes.resnet50()
n50 = param_shapes.numel(shapes)
gen = torch.Generator(device=dev).manual_seed(0)
g = [torch.randn(s, generator=gen, device=dev) for s in shapes]
p = [torch.randn(s, generator=gen, device=dev) * 0.02 for s in shapes]
m = [torch.zeros_like(t) for t in p]
v = [torch.zeros_like(t) for t in p]
Ls, La = mt.pack([g, g], 2), mt.pack([g, p, m, v], 4)
us = cold(lambda: lib.b200mt_scale(Ls.ref(), CHUNK, fp, 1.0, sp))
r["resnet_scale_us"], r["resnet_scale_frac"] = round(us, 2), round(8.0 * n50 / us / 1e3 / PEAK, 3)
us = cold(lambda: lib.b200mt_adam(La.ref(), CHUNK, fp, 1e-3, 0.9, 0.999, 1e-8, 7, 1, 1, 0.0, 1.0, sp))
r["resnet_adam_us"], r["resnet_adam_frac"] = round(us, 2), round(28.0 * n50 / us / 1e3 / PEAK, 3)


def step(k):
    f.zero_()
    mt.check(lib.b200mt_scale(Ls.ref(), CHUNK, fp, 1.0, sp))
    mt.check(lib.b200mt_adam(La.ref(), CHUNK, fp, 1e-3, 0.9, 0.999, 1e-8, k, 1, 1, 0.0, 1.0, sp))


for k in range(10):
    step(k + 1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for k in range(200):
    step(k + 11)
e1.record(stream)
torch.cuda.synchronize()
r["resnet_warm_step_us"] = round(e0.elapsed_time(e1) * 1e3 / 200, 2)
del g, p, m, v
if os.environ.get("SWEEP_ONLY") == "resnet":       # just the headline list (A/B of a change to the scale / Adam kernels)
    print(json.dumps(r))
    sys.exit(0)

# ---- mid-size lists
for tag, T, n in (("1000x64K", 1000, 65536), ("100x1M", 100, 1 << 20)):
    bx, xs = views(T, n)
    by, ys = views(T, n)
    bo, os_ = views(T, n)
    Lsc, Ln, Lax = mt.pack([xs, os_], 2), mt.pack([xs], 1), mt.pack([xs, ys, os_], 3)
    ret = torch.empty(1, device=dev)
    us = cold(lambda: lib.b200mt_scale(Lsc.ref(), CHUNK, fp, 0.5, sp))
    r[tag + "_scale_us"], r[tag + "_scale_frac"] = round(us, 2), round(8.0 * T * n / us / 1e3 / PEAK, 3)
    us = cold(lambda: lib.b200mt_l2norm(Ln.ref(), CHUNK, fp, ret.data_ptr(), None, 0, sp))
    r[tag + "_l2norm_us"], r[tag + "_l2norm_frac"] = round(us, 2), round(4.0 * T * n / us / 1e3 / PEAK, 3)
    us = cold(lambda: lib.b200mt_axpby(Lax.ref(), CHUNK, fp, 2.0, 0.5, -1, sp))
    r[tag + "_axpby_us"], r[tag + "_axpby_frac"] = round(us, 2), round(12.0 * T * n / us / 1e3 / PEAK, 3)
    del bx, by, bo, xs, ys, os_
    torch.cuda.empty_cache()

# ---- BERT-large gradient norm, fp16
shapes = param_shapes.bert_large()
nb = param_shapes.numel(shapes)
gb = torch.randn(nb, device=dev).half()
gl, off = [], 0
for s in shapes:
    k = 1
    for d in s:
        k *= d
    gl.append(gb[off:off + k].view(s)); off += k
Lg = mt.pack([gl], 1)
ret = torch.empty(1, device=dev)
us = cold(lambda: lib.b200mt_l2norm(Lg.ref(), CHUNK, fp, ret.data_ptr(), None, 0, sp))
r["bert_l2norm_fp16_us"], r["bert_l2norm_fp16_frac"] = round(us, 2), round(2.0 * nb / us / 1e3 / PEAK, 3)
del gb, gl

# ---- tiny-tensor regime through the Python entry points (host time dominates): wall time per call, queue kept short
bx, xs = views(10000, 1024)
by, ys = views(10000, 1024)
bo, os_ = views(10000, 1024)
for name, call in (("scale", lambda: amp_C.multi_tensor_scale(CHUNK, f, [xs, os_], 0.5)),
                   ("l2norm", lambda: amp_C.multi_tensor_l2norm(CHUNK, f, [xs], False)),
                   ("axpby", lambda: amp_C.multi_tensor_axpby(CHUNK, f, [xs, ys, os_], 2.0, 0.5, -1))):
    for _ in range(5):
        call()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        call()
    torch.cuda.synchronize()
    r["10kx1K_%s_us_per_call" % name] = round((time.perf_counter() - t0) / 50 * 1e6, 1)
# ... and with prepared list handles (amp_C.prepare): no walk, no image, no compare -- what is left is the launch
hs, hn, ha = amp_C.prepare([xs, os_]), amp_C.prepare([xs]), amp_C.prepare([xs, ys, os_])
for name, call in (("scale", lambda: amp_C.multi_tensor_scale(CHUNK, f, hs, 0.5)),
                   ("l2norm", lambda: amp_C.multi_tensor_l2norm(CHUNK, f, hn, False)),
                   ("axpby", lambda: amp_C.multi_tensor_axpby(CHUNK, f, ha, 2.0, 0.5, -1))):
    for _ in range(5):
        call()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        call()
    torch.cuda.synchronize()
    r["10kx1K_%s_prepared_us_per_call" % name] = round((time.perf_counter() - t0) / 50 * 1e6, 1)
print(json.dumps(r), flush=True)
