"""GPU: the REFERENCE's own Python package (apex.optimizers.FusedAdam / FusedLAMB / FusedSGD / FusedMixedPrecisionLamb,
apex.amp.scaler.LossScaler -- /root/reference/apex, unmodified, staged under oracle/_ref/pyref by oracle/build_ref.sh)
imported with this repo's `amp_C` shim first on sys.path, against the same classes running over the reference's own
kernels (oracle/_ref/amp_C_ref.so registered as `amp_C`).  This is the drop-in claim of INTEGRATION.md section 2:
"nothing in apex/ changes".  Two subprocesses (tests/_refpy_worker.py), so the two `amp_C` modules never meet.

Tolerances after 7 optimizer steps on parameters of order 1: Adam / SGD 1e-6 relative + 1e-6 absolute, LAMB (trust
ratios from norms) 1e-5 + 1e-6; fp16 model parameters one fp16 ulp; LossScaler outputs and flags bit-exact."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "_refpy_worker.py")


@pytest.fixture(scope="module")
def runs(tmp_path_factory):
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "pyref.zip")):
        pytest.skip("oracle/_ref/pyref not staged (oracle/build_ref.sh)")
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "amp_C_ref.so")):
        pytest.skip("oracle/_ref/amp_C_ref.so not built")
    d = tmp_path_factory.mktemp("refpy")
    out = {}
    for backend in ("ours", "ref"):
        path = str(d / (backend + ".pt"))
        env = dict(os.environ)
        env.pop("PYTHONPATH", None)
        r = subprocess.run([sys.executable, WORKER, backend, path], capture_output=True, text=True, timeout=900, env=env, cwd=str(d))
        assert r.returncode == 0, "%s worker failed:\n%s\n%s" % (backend, r.stdout[-2000:], r.stderr[-4000:])
        out[backend] = torch.load(path)
    return out


def _close(a, b, rtol, atol, what):
    assert a.shape == b.shape, what
    err = (a.double() - b.double()).abs()
    bad = err > (atol + rtol * b.double().abs())
    assert not bool(bad.any()), "%s: %d/%d outside rtol %g atol %g, worst %.3e" % (what, int(bad.sum()), err.numel(), rtol, atol, float(err.max()))


@pytest.mark.parametrize("tag,rtol", [("adam_w", 1e-6), ("adam_l2", 1e-6), ("sgd_mom", 1e-6), ("sgd_nesterov", 1e-6),
                                      ("lamb_nv", 1e-5), ("lamb_wd0", 1e-5), ("mplamb_fp32", 1e-5)])
@pytest.mark.parametrize("sizes", ["one", "multi"])
def test_reference_classes_give_the_same_parameters_over_both_back_ends(runs, tag, rtol, sizes):
    key = "%s/%s" % (tag, sizes)
    ours, ref = runs["ours"][key], runs["ref"][key]
    assert len(ours) == len(ref)
    moved = False
    for i, (a, b) in enumerate(zip(ours, ref)):
        _close(a, b, rtol, 1e-6, "%s param %d" % (key, i))
        moved = moved or float((b - b.mean()).abs().max()) > 0
    assert moved


def test_reference_mixed_precision_lamb_fp16_model(runs):
    ours, ref = runs["ours"], runs["ref"]
    assert ours["mplamb_fp16/multi/step"] == ref["mplamb_fp16/multi/step"] == 7
    for i, (a, b) in enumerate(zip(ours["mplamb_fp16/multi/masters"], ref["mplamb_fp16/multi/masters"])):
        _close(a, b, 1e-5, 1e-6, "fp32 master %d" % i)
    for i, (a, b) in enumerate(zip(ours["mplamb_fp16/multi"], ref["mplamb_fp16/multi"])):
        _close(a, b, 1e-3, 1e-6, "fp16 model param %d" % i)


def test_reference_loss_scaler_protocol(runs):
    ours, ref = runs["ours"], runs["ref"]
    for key in ("scaler/unscale", "scaler/unscale_with_stashed"):
        for a, b in zip(ours[key], ref[key]):
            assert torch.equal(a, b), key
    assert ours["scaler/flags"] == ref["scaler/flags"] == [False, False, True, 512.0]
