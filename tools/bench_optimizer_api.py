"""Probe: the optimizer classes themselves (apex.optimizers.FusedAdam / FusedLAMB / FusedSGD .step()) on the ResNet-50
list -- host time per step (list building + packing + launches) next to the device time, one B200."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import param_shapes
from apex.optimizers import FusedAdam, FusedLAMB, FusedSGD

dev = torch.device("cuda:0")
shapes = param_shapes.MODELS[os.environ.get("MODEL", "resnet50")]()
g = torch.Generator(device=dev).manual_seed(0)
N = sum(int(torch.Size(s).numel()) for s in shapes)
out = {"model": os.environ.get("MODEL", "resnet50"), "tensors": len(shapes), "params": N}
for name, make in (("FusedAdam", lambda ps: FusedAdam(ps, lr=1e-3)),
                   ("FusedLAMB", lambda ps: FusedLAMB(ps, lr=1e-3, weight_decay=0.01)),
                   ("FusedSGD", lambda ps: FusedSGD(ps, lr=1e-2, momentum=0.9))):
    params = [torch.nn.Parameter(torch.randn(s, generator=g, device=dev) * 0.02) for s in shapes]
    for p in params:
        p.grad = torch.randn_like(p) * 1e-3
    opt = make(params)
    for _ in range(5):
        opt.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 100
    e0.record()
    t0 = time.perf_counter()
    for _ in range(steps):
        opt.step()
    t1 = time.perf_counter()
    e1.record()
    torch.cuda.synchronize()
    out[name] = {"host_us_per_step": round((t1 - t0) / steps * 1e6, 1), "device_us_per_step": round(e0.elapsed_time(e1) / steps * 1e3, 1)}
print(json.dumps(out))
