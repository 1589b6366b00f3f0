"""Measures BASELINE.json configs 2, 3, 5 (and the unsharded Llama-3-8B mixed Adam) on ONE B200: every op of
this library next to the reference's own kernel (oracle/_ref/*.so, the reference sources compiled unmodified
for sm_100a) on identical inputs.  CUDA events around each call, L2 flushed (256 MiB write) before each timed
call, median of `reps`.  Prints JSON lines and writes a markdown table.  Not a product path: a measurement tool.

  python tools/bench_configs.py [--out profiles/configs_r01.md] [--skip-llama]
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import _b200mt as mt
import amp_C
import fused_adam_cuda
import param_shapes
from util import load_ref

PEAK = 6548.2
CHUNK = 2048 * 32
dev = torch.device("cuda:0")
flush_buf = None


def flush_l2():
    global flush_buf
    if flush_buf is None:
        flush_buf = torch.empty(64 << 20, dtype=torch.float32, device=dev)   # 256 MiB > 126 MB L2
    flush_buf.fill_(1.0)


def timeit(fn, reps=7, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = mt.lib.b200mt_launch_count()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
        launches = mt.lib.b200mt_launch_count() - l0
    return statistics.median(ts), launches


ROWS = []


def report(config, op, nelem, bytes_per_elem, ours_us, ours_launches, ref_us):
    gbs = nelem * bytes_per_elem / ours_us / 1e3
    row = {"config": config, "op": op, "elements": nelem, "algorithmic_B_per_elem": bytes_per_elem,
           "ours_us": round(ours_us, 1), "ours_launches": int(ours_launches), "ours_GBs": round(gbs, 1),
           "frac_of_6548": round(gbs / PEAK, 3), "params_per_s": nelem / ours_us * 1e6,
           "ref_us": None if ref_us is None else round(ref_us, 1),
           "speedup_vs_ref": None if ref_us is None else round(ref_us / ours_us, 2)}
    ROWS.append(row)
    print(json.dumps(row), flush=True)


def rand_list(shapes, dtype, scale=1.0, seed=0, positive=False):
    g = torch.Generator(device=dev).manual_seed(seed)
    out = []
    for s in shapes:
        t = torch.randn(s, generator=g, device=dev) * scale
        out.append((t.abs() if positive else t).to(dtype))
    return out


def flag():
    return torch.zeros(1, dtype=torch.int32, device=dev)


def config2(ref):
    shapes = param_shapes.resnet50()
    n = param_shapes.numel(shapes)
    g, p = rand_list(shapes, torch.float32, 65536.0, 0), rand_list(shapes, torch.float32, 0.02, 1)
    m, v = rand_list(shapes, torch.float32, 0.01, 2), rand_list(shapes, torch.float32, 0.001, 3, True)
    f = flag()
    o, l = timeit(lambda: amp_C.multi_tensor_scale(CHUNK, f, [g, g], 1.0))
    r, _ = timeit(lambda: ref.multi_tensor_scale(CHUNK, f, [g, g], 1.0)) if ref else (None, 0)
    report("2 resnet50", "multi_tensor_scale fp32 in-place", n, 8, o, l, r)
    args = (1e-3, 0.9, 0.999, 1e-8, 10, 1, 1, 0.0)
    o, l = timeit(lambda: amp_C.multi_tensor_adam(CHUNK, f, [g, p, m, v], *args))
    r, _ = timeit(lambda: ref.multi_tensor_adam(CHUNK, f, [g, p, m, v], *args)) if ref else (None, 0)
    report("2 resnet50", "multi_tensor_adam fp32", n, 28, o, l, r)
    # 10x replicated list (1610 tensors, 256 M params): working set >> L2
    shapes10 = shapes * 10
    g, p = rand_list(shapes10, torch.float32, 1.0, 0), rand_list(shapes10, torch.float32, 0.02, 1)
    m, v = rand_list(shapes10, torch.float32, 0.01, 2), rand_list(shapes10, torch.float32, 0.001, 3, True)
    o, l = timeit(lambda: amp_C.multi_tensor_adam(CHUNK, f, [g, p, m, v], *args))
    r, _ = timeit(lambda: ref.multi_tensor_adam(CHUNK, f, [g, p, m, v], *args)) if ref else (None, 0)
    report("2 resnet50 x10", "multi_tensor_adam fp32", n * 10, 28, o, l, r)


def config3(ref):
    shapes = param_shapes.bert_large()
    n = param_shapes.numel(shapes)
    f = flag()
    g16 = rand_list(shapes, torch.float16, 100.0, 0)
    p = rand_list(shapes, torch.float32, 0.02, 1)
    m, v = rand_list(shapes, torch.float32, 0.01, 2), rand_list(shapes, torch.float32, 0.001, 3, True)
    p16 = [x.half() for x in p]
    one = lambda val, dt=torch.float32: torch.tensor([val], dtype=dt, device=dev)
    lr, step, inv_scale, found_inf, max_norm = one(1e-3), one(5, torch.int32), one(1.0 / 128), one(0.0), one(128.0)
    keep = [x.clone() for x in g16]

    def restore():
        for a, b in zip(g16, keep):
            a.copy_(b)          # lamb overwrites g with the update: restore outside the timed call
        f.zero_()

    def mk(mod, nvlamb):
        def run():
            gn, _ = mod.multi_tensor_l2norm_mp(CHUNK, f, [g16], False)
            mod.multi_tensor_lamb_mp(CHUNK, f, [g16, p, m, v, p16], lr, 0.9, 0.999, 1e-6, step, 1, 0.01, 1, 1, gn, max_norm,
                                     nvlamb, found_inf, inv_scale)
        return run

    def timed_with_restore(fn, reps=5):
        ts, launches = [], 0
        for i in range(reps + 1):
            restore(); flush_l2()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = mt.lib.b200mt_launch_count()
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            launches = mt.lib.b200mt_launch_count() - l0
            if i:
                ts.append(e0.elapsed_time(e1) * 1e3)
        return statistics.median(ts), launches

    for nvlamb in (False, True):
        o, l = timed_with_restore(mk(amp_C, nvlamb))
        r, _ = timed_with_restore(mk(ref, nvlamb)) if ref else (None, 0)
        report("3 bert_large", "l2norm_mp + lamb_mp (g16,p32,m32,v32,p16) nvlamb=%s" % nvlamb, n, 44, o, l, r)
    # all-fp32 variant
    g32 = [x.float() for x in keep]
    keep32 = [x.clone() for x in g32]

    def restore32():
        for a, b in zip(g32, keep32):
            a.copy_(b)
        f.zero_()

    def mk32(mod):
        def run():
            gn, _ = mod.multi_tensor_l2norm(CHUNK, f, [g32], False)
            mod.multi_tensor_lamb(CHUNK, f, [g32, p, m, v], 1e-3, 0.9, 0.999, 1e-6, 5, 1, 0.01, 1, 1, gn, 1.0, False)
        return run

    def timed32(fn, reps=5):
        ts, launches = [], 0
        for i in range(reps + 1):
            restore32(); flush_l2()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = mt.lib.b200mt_launch_count()
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            launches = mt.lib.b200mt_launch_count() - l0
            if i:
                ts.append(e0.elapsed_time(e1) * 1e3)
        return statistics.median(ts), launches

    o, l = timed32(mk32(amp_C))
    r, _ = timed32(mk32(ref)) if ref else (None, 0)
    report("3' bert_large", "l2norm + lamb all-fp32", n, 52, o, l, r)


def config5(ref):
    f = flag()
    for T, n in [(10000, 1024), (1000, 65536), (100, 1 << 20), (16, 1 << 26), (4, 1 << 28)]:
        big_x = torch.randn(T * n, device=dev)
        big_y = torch.randn(T * n, device=dev)
        big_o = torch.empty(T * n, device=dev)
        xs, ys, os_ = list(big_x.view(T, n).unbind(0)), list(big_y.view(T, n).unbind(0)), list(big_o.view(T, n).unbind(0))
        tag = "5 %dx%d" % (T, n)
        o, l = timeit(lambda: amp_C.multi_tensor_scale(CHUNK, f, [xs, os_], 0.5))
        r, _ = timeit(lambda: ref.multi_tensor_scale(CHUNK, f, [xs, os_], 0.5)) if ref else (None, 0)
        report(tag, "scale fp32->fp32", T * n, 8, o, l, r)
        o, l = timeit(lambda: amp_C.multi_tensor_l2norm(CHUNK, f, [xs], False))
        r, _ = timeit(lambda: ref.multi_tensor_l2norm(CHUNK, f, [xs], False)) if ref else (None, 0)
        report(tag, "l2norm global", T * n, 4, o, l, r)
        o, l = timeit(lambda: amp_C.multi_tensor_l2norm(CHUNK, f, [xs], True))
        r, _ = timeit(lambda: ref.multi_tensor_l2norm(CHUNK, f, [xs], True)) if ref else (None, 0)
        report(tag, "l2norm per-tensor", T * n, 4, o, l, r)
        o, l = timeit(lambda: amp_C.multi_tensor_axpby(CHUNK, f, [xs, ys, os_], 2.0, 0.5, -1))
        r, _ = timeit(lambda: ref.multi_tensor_axpby(CHUNK, f, [xs, ys, os_], 2.0, 0.5, -1)) if ref else (None, 0)
        report(tag, "axpby fp32", T * n, 12, o, l, r)
        xh = big_x.half()
        xhs = list(xh.view(T, n).unbind(0))
        o, l = timeit(lambda: amp_C.multi_tensor_scale(CHUNK, f, [xhs, os_], 0.5))
        r, _ = timeit(lambda: ref.multi_tensor_scale(CHUNK, f, [xhs, os_], 0.5)) if ref else (None, 0)
        report(tag, "scale fp16->fp32", T * n, 6, o, l, r)
        del big_x, big_y, big_o, xs, ys, os_, xh, xhs
        torch.cuda.empty_cache()
    # odd sizes / misaligned views: the reference's scalar-fallback regime
    base = torch.randn(64 * (2048 * 32 + 1) + 8, device=dev)
    outb = torch.empty_like(base)
    n = 2048 * 32 + 1
    xs = [base[1 + i * n:1 + (i + 1) * n] for i in range(64)]
    os_ = [outb[1 + i * n:1 + (i + 1) * n] for i in range(64)]
    o, l = timeit(lambda: amp_C.multi_tensor_scale(CHUNK, f, [xs, os_], 0.5))
    r, _ = timeit(lambda: ref.multi_tensor_scale(CHUNK, f, [xs, os_], 0.5)) if ref else (None, 0)
    report("5 64x65537 misaligned views", "scale fp32->fp32", 64 * n, 8, o, l, r)


def llama_unsharded():
    refc = load_ref("fused_adam_cuda_ref")
    shapes = param_shapes.llama3_8b()
    n = param_shapes.numel(shapes)
    f = flag()
    mk = lambda dt, val: [torch.full(s, val, dtype=dt, device=dev) for s in shapes]
    p, m, v = mk(torch.float32, 0.01), mk(torch.float32, 0.001), mk(torch.float32, 0.0001)
    g, pc = mk(torch.float16, 0.5), mk(torch.float16, 0.0)
    args = (1e-4, 0.9, 0.95, 1e-8, 1.0, 3, 1, 1, 0.1)
    o, l = timeit(lambda: fused_adam_cuda.adam_mt(CHUNK, f, [p, m, v, g, pc], *args), reps=3, warm=1)
    r, _ = timeit(lambda: refc.adam_mt(CHUNK, f, [p, m, v, g, pc], *args), reps=3, warm=1) if refc else (None, 0)
    report("4' llama3_8b unsharded", "contrib adam_mt (p32,m32,v32,g16,p16)", n, 28, o, l, r)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.md"))
    ap.add_argument("--skip-llama", action="store_true")
    a = ap.parse_args()
    ref = load_ref("amp_C_ref")
    config2(ref)
    config3(ref)
    config5(ref)
    if not a.skip_llama:
        llama_unsharded()
    with open(a.out, "w") as fh:
        fh.write("| config | op | elements | alg. B/elem | ours µs | ours launches | ours GB/s | frac of 6548 GB/s | ref µs | speed-up vs ref |\n")
        fh.write("|---|---|---:|---:|---:|---:|---:|---:|---:|---:|\n")
        for r in ROWS:
            fh.write("| %s | %s | %d | %d | %.1f | %d | %.1f | %.3f | %s | %s |\n" % (
                r["config"], r["op"], r["elements"], r["algorithmic_B_per_elem"], r["ours_us"], r["ours_launches"], r["ours_GBs"],
                r["frac_of_6548"], r["ref_us"], r["speedup_vs_ref"]))
    print("wrote", a.out)


if __name__ == "__main__":
    main()
