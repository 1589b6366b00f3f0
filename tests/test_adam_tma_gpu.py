"""GPU: the opt-in bulk-copy (TMA) staging variant of the fp32 multi_tensor_adam pass (csrc/ops_adam_tma.cu,
B200MT_ADAM_TMA=1; reference op: csrc/multi_tensor_adam.cu:23-127).  The switch is read once per process, so the
check runs tools/adam_tma_probe.py in a child process: ragged sizes (tails of 1-3 elements, tensors smaller than a
quad, several tiles) against a float64 evaluation of the rule, both decay modes, tolerance 1e-6 absolute."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("tile", ["2048", "4096"])
def test_bulk_copy_adam_matches_the_rule(tile):
    env = dict(os.environ, B200MT_ADAM_TMA="1", B200MT_ADAM_TMA_TILE=tile)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "adam_tma_probe.py")], env=env,
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["tma"] == "1" and d["max_abs_err_mode0"] < 1e-6 and d["max_abs_err_mode1"] < 1e-6
    assert d["resnet50"]["launches"] == 100          # one launch per call on this path too
