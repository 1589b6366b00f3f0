"""Probe (not product): per-kernel durations (library launch trace) of the light ops on the shapes the round-1 verdict
flagged, L2 flushed before every call; the same script is the workload of the ncu captures under profiles/
(`ncu --set full -k regex:mt_stream_kernel|norm_cleanup ... python tools/prof_light.py --ncu`)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch

import _b200mt as mt
import param_shapes

dev = torch.device("cuda:0")
lib, CHUNK, PEAK = mt.lib, 65536, 6548.2
stream = torch.cuda.current_stream(dev)
sp = stream.cuda_stream
f = torch.zeros(1, dtype=torch.int32, device=dev)
fp = f.data_ptr()
flush = torch.empty(128 << 20, device=dev)
ncu = "--ncu" in sys.argv

shapes = param_shapes.resnet50()
n50 = param_shapes.numel(shapes)
g = [torch.randn(s, device=dev) for s in shapes]
Ls = mt.pack([g, g], 2)
big = torch.randn(1000 * 65536, device=dev)
out = torch.empty_like(big)
xs, os_ = list(big.view(1000, 65536).unbind(0)), list(out.view(1000, 65536).unbind(0))
Ln, Lsc = mt.pack([xs], 1), mt.pack([xs, os_], 2)
ret = torch.empty(1, device=dev)
calls = [("resnet50 scale fp32 in place", 8.0 * n50, lambda: lib.b200mt_scale(Ls.ref(), CHUNK, fp, 1.0, sp)),
         ("1000x64K l2norm fp32", 4.0 * 65536000, lambda: lib.b200mt_l2norm(Ln.ref(), CHUNK, fp, ret.data_ptr(), None, 0, sp)),
         ("1000x64K scale fp32", 8.0 * 65536000, lambda: lib.b200mt_scale(Lsc.ref(), CHUNK, fp, 0.5, sp))]
for name, nbytes, call in calls:
    mt.check(call()); mt.check(call())
    torch.cuda.synchronize()
    if ncu:
        flush.fill_(1.0)
        mt.check(call())
        torch.cuda.synchronize()
        continue
    per = {}
    for k in range(7):
        flush.fill_(float(k))
        with mt.trace(8) as tr:
            mt.check(call())
            torch.cuda.synchronize()
        for kn, ms in tr.launches:
            per.setdefault(kn.replace("b200mt::", ""), []).append(ms * 1e3)
    row = {"case": name, "kernels": {k: round(sorted(v)[len(v) // 2], 2) for k, v in per.items()}}
    tot = sum(row["kernels"].values())
    row["sum_us"], row["frac_of_6548"] = round(tot, 2), round(nbytes / tot / 1e3 / PEAK, 3)
    print(json.dumps(row), flush=True)
