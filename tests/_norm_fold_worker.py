"""Subprocess body of tests/test_norm_fold_gpu.py: norm-type ops on seeded inputs, results saved for a bitwise comparison
between the two-kernel form (default) and the single-launch form (B200MT_NORM_FOLD=1, read once per process)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "rocm-apex_b200"), ROOT]
import torch

import amp_C
import _b200mt as mt


def main():
    out = sys.argv[1]
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(77)
    sizes = [1, 0, 3, 555, 2048 * 32 + 1, 33333, 0, 8192, 16, 2048 * 32 * 5 - 1, 300001]
    res = {}
    for dt in (torch.float32, torch.float16, torch.bfloat16):
        xs = [(torch.randn(n, generator=g, device=dev) * 3).to(dt) for n in sizes]
        for chunk in (65536, 333):
            f = torch.zeros(1, dtype=torch.int32, device=dev)
            n0 = mt.lib.b200mt_launch_count()
            norm, per = amp_C.multi_tensor_l2norm(chunk, f, [xs], True)
            res["l2norm/%s/%d" % (dt, chunk)] = (norm.cpu(), per.cpu(), int(f), mt.lib.b200mt_launch_count() - n0)
        # a chunk whose sum of squares overflows although every element is finite -> flag
        big = [torch.full((70000,), 3e19 if dt != torch.float16 else 60000.0, device=dev).to(dt), xs[3]]
        f = torch.zeros(1, dtype=torch.int32, device=dev)
        norm, per = amp_C.multi_tensor_l2norm(65536, f, [big], True)
        res["overflow/%s" % dt] = (norm.cpu(), per.cpu(), int(f))
        outs = [torch.empty(n, device=dev) for n in sizes]
        norm, per = amp_C.multi_tensor_l2norm_scale(65536, f, [xs, outs], 0.5, True)
        res["l2norm_scale/%s" % dt] = (norm.cpu(), per.cpu(), [o.cpu() for o in outs])
    # gated variant with the flag already set: nothing may be written
    f = torch.ones(1, dtype=torch.int32, device=dev)
    xs = [torch.randn(n, generator=g, device=dev) for n in sizes]
    norm, per = amp_C.multi_tensor_l2norm_mp(65536, f, [xs], True)
    res["gated_skipped"] = int(f)
    # LAMB (two sets of partials, per-tensor trust ratios) and mixed-precision LAMB
    sz = [n for n in sizes if n]
    mk = lambda s, pos=False: [((torch.randn(n, generator=g, device=dev) * s).abs() if pos else torch.randn(n, generator=g, device=dev) * s) for n in sz]
    lists = [mk(1.0), mk(0.02), mk(0.01), mk(0.001, True)]
    f = torch.zeros(1, dtype=torch.int32, device=dev)
    gn, _ = amp_C.multi_tensor_l2norm(65536, f, [lists[0]], False)
    n0 = mt.lib.b200mt_launch_count()
    amp_C.multi_tensor_lamb(65536, f, lists, 1e-2, 0.9, 0.999, 1e-6, 3, 1, 0.01, 1, 1, gn, 1.0, True)
    res["lamb"] = ([[t.cpu() for t in l] for l in lists], int(f), mt.lib.b200mt_launch_count() - n0)
    g16 = [t.half() for t in mk(64.0)]
    lists = [g16, mk(0.02), mk(0.01), mk(0.001, True)]
    lists.append([t.half() for t in lists[1]])
    one = lambda v, d=torch.float32: torch.tensor([v], dtype=d, device=dev)
    gn, _ = amp_C.multi_tensor_l2norm_mp(65536, f, [g16], False)
    amp_C.multi_tensor_lamb_mp(65536, f, lists, one(2e-3), 0.9, 0.999, 1e-6, one(2, torch.int32), 1, 0.01, 1, 1, gn, one(64.0), False,
                               one(0.0), one(1.0 / 64))
    res["lamb_mp"] = ([[t.cpu() for t in l] for l in lists], int(f))
    # back-to-back calls on one stream reuse the tickets: a second, different list right away
    ys = [torch.randn(5000, generator=g, device=dev) for _ in range(40)]
    a = amp_C.multi_tensor_l2norm(65536, f, [ys], True)
    b = amp_C.multi_tensor_l2norm(65536, f, [ys[:7]], True)
    res["back_to_back"] = (a[0].cpu(), a[1].cpu(), b[0].cpu(), b[1].cpu())
    torch.cuda.synchronize()
    torch.save(res, out)


if __name__ == "__main__":
    main()
