"""GPU: the REFERENCE's own unit tests for this path, run unmodified (oracle/_ref/pyref_tests, staged from
/root/reference/tests/L0/run_amp and run_optimizers by oracle/build_ref.sh) with the reference's own Python package
over this repo's `amp_C` shim -- and, for comparison, over the reference's own kernels (amp_C_ref.so).

The reference targets an older torch: some of its tests fail on torch 2.11 for reasons that have nothing to do with the
kernels (removed torch APIs in the amp patch lists, ...).  The contract checked here is therefore differential: every
reference test that passes over the reference's own kernels must pass over this repo's shim, and the tests that pin the
kernels' known answers (multi_tensor_scale / axpby / l2norm, the fused optimizers, LAMB, FusedSGD + amp) must be among
those that pass.  The outcome table of both runs is written to gpurun_out/reference_suite_r02.json."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUNNER = os.path.join(ROOT, "tests", "_reftests_runner.py")
FILES = ["run_amp/test_multi_tensor_scale.py", "run_amp/test_multi_tensor_axpby.py", "run_amp/test_multi_tensor_l2norm.py",
         "run_amp/test_fused_sgd.py", "run_amp/test_multiple_models_optimizers_losses.py", "run_amp/test_checkpointing.py",
         "run_amp/test_add_param_group.py", "run_optimizers/test_fused_optimizer.py", "run_optimizers/test_lamb.py",
         "run_optimizers/test_fused_novograd.py"]


@pytest.fixture(scope="module")
def outcomes(tmp_path_factory):
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "pyref_tests.zip")):
        pytest.skip("oracle/_ref/pyref_tests not staged (oracle/build_ref.sh)")
    d = tmp_path_factory.mktemp("refsuite")
    res = {}
    for backend in ("ours", "ref"):
        path = str(d / (backend + ".json"))
        env = dict(os.environ)
        env.pop("PYTHONPATH", None)
        r = subprocess.run([sys.executable, RUNNER, backend, path] + FILES, capture_output=True, text=True, timeout=1500, env=env,
                           cwd=str(d))
        assert os.path.exists(path), "%s runner died:\n%s\n%s" % (backend, r.stdout[-2000:], r.stderr[-4000:])
        res[backend] = json.load(open(path))
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "reference_suite_r02.json"), "w") as fh:
        json.dump(res, fh, indent=1)
    return res


def test_every_reference_test_that_passes_on_the_reference_kernels_passes_on_the_shim(outcomes):
    ours, ref = outcomes["ours"], outcomes["ref"]
    assert set(ours) == set(ref) and len(ours) > 20
    regress = {k: ours[k] for k in ref if ref[k] == "ok" and ours[k] != "ok"}
    assert not regress, "reference tests that pass on amp_C_ref but not on the shim:\n" + json.dumps(regress, indent=1)[:6000]
    print("reference tests: %d collected, %d ok over the shim, %d ok over amp_C_ref" % (
        len(ours), sum(v == "ok" for v in ours.values()), sum(v == "ok" for v in ref.values())))


@pytest.mark.parametrize("needle", ["test_multi_tensor_scale", "test_multi_tensor_axpby", "test_multi_tensor_l2norm",
                                    "test_fused_optimizer", "test_lamb", "test_fused_sgd"])
def test_kernel_pinning_reference_tests_pass_on_the_shim(outcomes, needle):
    mine = {k: v for k, v in outcomes["ours"].items() if needle in k}
    assert mine, "no reference test matched %s" % needle
    passed = [k for k, v in mine.items() if v == "ok"]
    assert passed, "none of the reference's %s tests passed on the shim: %s" % (needle, json.dumps(mine, indent=1)[:3000])
