"""Golden fixtures recorded from the reference's own Python (tests/golden/make_golden.py imports
apex/amp/scaler.py from the reference tree): the oracle must reproduce them on CPU, the CUDA kernels on GPU.

The reference's python fallback decides overflow from the SUM of the tensor (apex/amp/scaler.py:9-14) and
returns before writing anything; the fused kernel tests every element and still writes the output.  On these
fixtures both agree on the flag; values are compared only where the reference wrote them."""
import os

import pytest
import torch

from oracle import mt_oracle as O

GOLD = torch.load(os.path.join(os.path.dirname(__file__), "golden", "ref_python_fallbacks.pt"))


def _tol(dt):
    # fp16 outputs: the python fallback copies (rounds) into the fp16 master BEFORE multiplying, the fused
    # kernel multiplies in fp32 and rounds once (csrc/multi_tensor_scale_kernel.cu:66-70): two half-ulps apart
    return {torch.float32: 2e-7, torch.float16: 2e-3, torch.bfloat16: 8e-3}[dt]


def test_oracle_reproduces_reference_python_fallbacks():
    n = 0
    for c in GOLD["cases"]:
        flag = [0]
        if c["op"] == "scale":
            out = torch.zeros_like(c["master_out"])
            O.multi_tensor_scale(flag, [[c["model"]], [out]], c["scale"])
        else:
            out = torch.zeros_like(c["master_out"])
            O.multi_tensor_axpby(flag, [[c["model"]], [c["stashed"]], [out]], c["a"], c["b"], 0)
        assert bool(flag[0]) == c["overflow"]
        if not c["overflow"]:
            assert torch.allclose(out.float(), c["master_out"].float(), rtol=_tol(out.dtype), atol=1e-30)
            n += 1
    assert n >= 30


def test_oracle_port_of_fallbacks_is_the_reference_function():
    # oracle.scale_check_overflow_python is what bench.py times as the CPU baseline
    for c in GOLD["cases"]:
        if c["op"] != "scale":
            continue
        out = torch.zeros_like(c["master_out"])
        ov = O.scale_check_overflow_python(c["model"], out, c["scale"], check_overflow=True)
        assert ov == c["overflow"]
        if not ov:
            assert torch.equal(out, c["master_out"])


@pytest.mark.gpu
def test_kernels_reproduce_reference_python_fallbacks():
    import amp_C
    for c in GOLD["cases"]:
        flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        out = torch.zeros_like(c["master_out"]).cuda()
        if c["op"] == "scale":
            amp_C.multi_tensor_scale(65536, flag, [[c["model"].cuda()], [out]], c["scale"])
        else:
            amp_C.multi_tensor_axpby(65536, flag, [[c["model"].cuda()], [c["stashed"].cuda()], [out]], c["a"], c["b"], 0)
        assert bool(int(flag)) == c["overflow"]
        if not c["overflow"]:
            want = c["master_out"].float()
            # a*x + b*y is one FMA on the GPU and two roundings in the reference's Python: an ulp of the
            # operands apart, which is large relative to results that cancel -> floor scaled by max|want|
            atol = 1e-30 if c["op"] == "scale" else _tol(out.dtype) * float(want.abs().max())
            assert torch.allclose(out.cpu().float(), want, rtol=_tol(out.dtype), atol=atol)


@pytest.mark.gpu
def test_kernels_reproduce_reference_known_answers():
    import amp_C
    k = GOLD["kats"]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    x = [torch.full((n,), k["scale"]["value"], device="cuda") for n in (777, 65537)]
    amp_C.multi_tensor_scale(65536, f, [x, x], k["scale"]["scale"])
    assert all(torch.equal(t, torch.full_like(t, k["scale"]["expect"])) for t in x) and int(f) == 0
    a = k["axpby"]
    xs, ys = [torch.full((4099,), a["x"], device="cuda")], [torch.full((4099,), a["y"], device="cuda")]
    out = [torch.zeros(4099, device="cuda")]
    amp_C.multi_tensor_axpby(65536, f, [xs, ys, out], a["a"], a["b"], -1)
    assert torch.equal(out[0], torch.full_like(out[0], a["expect"]))
    for sizea, sizeb, rep in k["l2norm"]["sizes"]:
        ts = [torch.full((sizea,), k["l2norm"]["value"], device="cuda"), torch.full((sizeb,), k["l2norm"]["value"], device="cuda")] * rep
        n, per = amp_C.multi_tensor_l2norm(65536, f, [ts], True)
        want = k["l2norm"]["value"] * ((sizea + sizeb) * rep) ** 0.5
        assert abs(float(n) - want) <= 1e-5 * want
        assert abs(float(per[0]) - k["l2norm"]["value"] * sizea ** 0.5) <= 1e-5 * float(per[0])
