"""Generates tests/golden/ref_python_fallbacks.pt by IMPORTING THE REFERENCE'S OWN PYTHON
(/root/reference, unmodified, with a 3-line shim for the removed torch._six module) and recording
what its python-only build computes for the amp unscale path on seeded inputs:

  apex.amp.scaler.scale_check_overflow_python   (apex/amp/scaler.py:6-23)
  apex.amp.scaler.axpby_check_overflow_python   (apex/amp/scaler.py:25-40)

plus the known answers the reference's kernel tests assert (tests/L0/run_amp/test_multi_tensor_scale.py,
test_multi_tensor_axpby.py, test_multi_tensor_l2norm.py).  Run here (the reference does not travel to the
GPU box):  python tests/golden/make_golden.py
"""
import collections.abc
import os
import sys
import types

import torch

REF = os.environ.get("REF", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_python_fallbacks.pt")


def import_reference_scaler():
    six = types.ModuleType("torch._six")
    six.string_classes = (str, bytes)
    six.container_abcs = collections.abc
    sys.modules["torch._six"] = six
    sys.path.insert(0, REF)
    from apex.amp import scaler      # the reference's module, python-only build (no amp_C on this path)
    from apex.multi_tensor_apply import multi_tensor_applier
    assert not multi_tensor_applier.available, "expected the python-only build of the reference"
    return scaler


def main():
    scaler = import_reference_scaler()
    g = torch.Generator().manual_seed(20240607)
    cases = []
    for in_dt, out_dt in [(torch.float32, torch.float32), (torch.float16, torch.float32), (torch.bfloat16, torch.float32),
                          (torch.float32, torch.float16)]:
        for n in (1, 555, 2051):
            for scale, poison in [(1.0 / 65536.0, None), (0.5, None), (1.0 / 128.0, "inf"), (1.0 / 128.0, "nan")]:
                model = (torch.randn(n, generator=g) * 1000.0).to(in_dt)
                if poison:
                    model[n // 2] = float(poison)
                master = torch.zeros(n, dtype=out_dt)
                overflow = scaler.scale_check_overflow_python(model, master, scale, check_overflow=True)
                cases.append({"op": "scale", "model": model, "scale": scale, "master_out": master.clone(),
                              "overflow": bool(overflow)})
    for dt in (torch.float32, torch.float16):
        for n in (7, 2051):
            for a, b in [(2.0, 8.0), (1.0 / 64.0, 1.0), (0.37, -1.5)]:
                model = (torch.randn(n, generator=g) * 10.0).to(dt)
                stashed = torch.randn(n, generator=g)
                master = torch.nn.Parameter(torch.zeros(n))
                overflow = scaler.axpby_check_overflow_python(model, stashed, master, a, b, check_overflow=True)
                cases.append({"op": "axpby", "model": model, "stashed": stashed, "a": a, "b": b,
                              "master_out": master.data.clone(), "overflow": bool(overflow)})
    kats = {
        "scale": {"value": 4.0, "scale": 0.25, "expect": 1.0},                 # test_multi_tensor_scale.py:36-56
        "axpby": {"x": 4.0, "y": 16.0, "a": 2.0, "b": 8.0, "expect": 136.0},   # test_multi_tensor_axpby.py:34-39
        "l2norm": {"value": 4.0, "sizes": [[7777 * 77, 555, 1], [2048 * 32 + 1, 333, 55], [33333, 555, 3]]},   # test_multi_tensor_l2norm.py:35-56
    }
    torch.save({"cases": cases, "kats": kats, "reference": "abhinavvishnu/rocm-apex apex/amp/scaler.py:6-40"}, OUT)
    print("wrote", OUT, os.path.getsize(OUT), "bytes,", len(cases), "cases")


if __name__ == "__main__":
    main()
