"""GPU: the opt-in single-launch form of the norm-type ops (B200MT_NORM_FOLD=1: the second pass runs inside the streaming
kernel, done by the CTA that finishes each tensor last -- csrc/mt_norm.cuh) must give BITWISE the results of the default
two-kernel form: same partial slots, same summation order, same flag.  Two subprocesses, since the switch is read once."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _same(a, b, path="res"):
    if isinstance(a, torch.Tensor):
        assert torch.equal(a, b) or (torch.equal(a.isnan(), b.isnan()) and torch.equal(a.nan_to_num(), b.nan_to_num())), path
    elif isinstance(a, (list, tuple)):
        assert len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, "%s[%d]" % (path, i))
    else:
        assert a == b, (path, a, b)


def test_folded_norm_form_is_bitwise_identical(tmp_path):
    res = {}
    for fold in ("0", "1"):
        out = str(tmp_path / ("fold%s.pt" % fold))
        env = dict(os.environ, B200MT_NORM_FOLD=fold)
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "_norm_fold_worker.py"), out], capture_output=True, text=True,
                           timeout=600, env=env)
        assert r.returncode == 0, r.stderr[-3000:]
        res[fold] = torch.load(out)
    assert set(res["0"]) == set(res["1"])
    for key in res["0"]:
        a, b = res["0"][key], res["1"][key]
        if key.startswith("l2norm/"):          # last field: kernels launched -- 2 in the default form, 1 folded
            assert a[3] == 2 and b[3] == 1, (key, a[3], b[3])
            a, b = a[:3], b[:3]
        if key == "lamb":
            assert a[2] == 3 and b[2] == 2, (a[2], b[2])
            a, b = a[:2], b[:2]
        _same(a, b, key)
    assert res["0"]["gated_skipped"] == 1
    assert all(res["0"]["overflow/%s" % dt][2] == 1 for dt in (torch.float32, torch.bfloat16))
