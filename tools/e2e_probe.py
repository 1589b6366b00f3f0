"""Probe: where the end-to-end ResNet-50 step (gradients arriving from pinned host memory) spends its time, and
what splitting the copy into segments that overlap with the unscale + Adam launches buys.  One B200."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import amp_C, param_shapes

dev = torch.device("cuda:0")
shapes = param_shapes.resnet50()
numels = [int(torch.Size(s).numel()) for s in shapes]
N = sum(numels)
host = torch.empty(N, dtype=torch.float32).pin_memory()
host.normal_(0, 1).mul_(65536.0)
flat = torch.empty(N, dtype=torch.float32, device=dev)
g = torch.Generator(device=dev).manual_seed(0)
params = [torch.randn(s, generator=g, device=dev) * 0.02 for s in shapes]
m = [torch.zeros_like(p) for p in params]
v = [torch.zeros_like(p) for p in params]
views, off = [], 0
for s, n in zip(shapes, numels):
    views.append(flat[off:off + n].view(s)); off += n
flag = torch.zeros(1, dtype=torch.int32, device=dev)
cur = torch.cuda.current_stream(dev)
copy_stream = torch.cuda.Stream(device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(fn, reps=20):
    for k in range(3):
        fn(k + 1)
    torch.cuda.synchronize()
    e0.record(cur)
    for k in range(reps):
        fn(k + 4)
    e1.record(cur)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def copy_only(k):
    flat.copy_(host, non_blocking=True)


def plain(k):
    flat.copy_(host, non_blocking=True)
    flag.zero_()
    amp_C.multi_tensor_scale(65536, flag, [views, views], 1.0 / 65536)
    amp_C.multi_tensor_adam(65536, flag, [views, params, m, v], 1e-3, 0.9, 0.999, 1e-8, k, 1, 1, 0.0)
    return flag.item()


def segmented(nseg):
    # segment boundaries on tensor boundaries, roughly equal bytes
    bounds, acc, target = [0], 0, N / nseg
    for i, n in enumerate(numels):
        acc += n
        if acc >= target * len(bounds) and len(bounds) < nseg:
            bounds.append(i + 1)
    bounds.append(len(numels))
    offs = [0]
    for n in numels:
        offs.append(offs[-1] + n)
    evs = [torch.cuda.Event() for _ in range(nseg)]

    def fn(k):
        flag.zero_()
        copy_stream.wait_stream(cur)            # the previous step has finished with `flat`
        with torch.cuda.stream(copy_stream):
            for s in range(nseg):
                a, b = offs[bounds[s]], offs[bounds[s + 1]]
                flat[a:b].copy_(host[a:b], non_blocking=True)
                evs[s].record(copy_stream)
        for s in range(nseg):
            cur.wait_event(evs[s])
            sl = slice(bounds[s], bounds[s + 1])
            amp_C.multi_tensor_scale(65536, flag, [views[sl], views[sl]], 1.0 / 65536)
            amp_C.multi_tensor_adam(65536, flag, [views[sl], params[sl], m[sl], v[sl]], 1e-3, 0.9, 0.999, 1e-8, k, 1, 1, 0.0)
        return flag.item()
    return fn


r = {"params": N, "copy_only_ms": timed(copy_only), "plain_ms": timed(plain)}
r["copy_GBs"] = 4 * N / r["copy_only_ms"] / 1e6
for nseg in (2, 4, 8):
    r["segmented_%d_ms" % nseg] = timed(segmented(nseg))
print(json.dumps(r))
