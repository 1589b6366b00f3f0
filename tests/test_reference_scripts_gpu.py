"""GPU: the REFERENCE's own distributed test SCRIPTS, run unmodified against THIS repo's `apex` package.  First and foremost
its script for the sharded optimizer (tests/L0/run_optimizers/test_dist_adam.py, staged unmodified as
bytecode in oracle/_ref/pyref_tests.zip by oracle/build_ref.sh) run against THIS repo's `apex` package: its
`DistributedFusedAdam` (constructor keywords, `set_global_scale`, `_init_everything`, `complete_reductions`, `step`), its `amp`
(O2, dynamic loss scale) and its `FusedAdam`, next to torch's DistributedDataParallel over NCCL.

The script trains a small half-precision model with the sharded optimizer and an fp32-master twin with FusedAdam, asserts that
inputs, outputs and input gradients agree every step, and compares the new weights with its own tolerances; on a weight mismatch
it prints the offending parameter ("Rank: r, Param: i") and leaves with status 0, so the output is checked for that line too.

The reference's own DistributedFusedAdam CLASS cannot be the one under test on this torch: it asserts a `no_copy` keyword of
`torch.distributed.reduce_scatter` that only its authors' c10d fork has (apex/contrib/optimizers/distributed_fused_adam.py:371).

Also here, two ranks each: tests/distributed/DDP/ddp_race_condition_test.py (the DDP gradient-bucket path) and
tests/distributed/amp_master_params/{amp_master_params,compare}.py (amp O2 master weights under apex DDP)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TESTS_ZIP = os.path.join(ROOT, "oracle", "_ref", "pyref_tests.zip")
STEPS = 8

BODY = r"""
import runpy, sys
sys.path[:0] = [%(pkg)r, %(tests)r]
import apex, os
assert os.path.realpath(apex.__file__).startswith(os.path.realpath(%(pkg)r)), apex.__file__       # this repo's package
sys.argv = ["test_dist_adam.py", "--local_rank", os.environ["LOCAL_RANK"], "--steps", "%(steps)d", "--dim", "%(dim)d",
            "--layers", "%(layers)d"] + %(extra)r
runpy.run_module("test_dist_adam", run_name="__main__")
"""


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, dim, layers, extra=()):
    if not os.path.exists(TESTS_ZIP):
        pytest.skip("oracle/_ref/pyref_tests.zip not staged (oracle/build_ref.sh)")
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    port = _free_port()
    body = BODY % dict(pkg=os.path.join(ROOT, "rocm-apex_b200"), tests=os.path.join(TESTS_ZIP, "run_optimizers"), steps=STEPS,
                       dim=dim, layers=layers, extra=list(extra))
    procs = []
    for r in range(world):
        env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(r), LOCAL_RANK=str(r),
                   WORLD_SIZE=str(world))
        env.pop("PYTHONPATH", None)
        procs.append(subprocess.Popen([sys.executable, "-c", body], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                      text=True, cwd=ROOT))
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, o, e))
    for r, (rc, o, e) in enumerate(outs):
        assert rc == 0, "rank %d left with %s\n%s\n%s" % (r, rc, o[-3000:], e[-4000:])
        assert o.count("Checking new weights") == STEPS, "rank %d stopped early:\n%s\n%s" % (r, o[-3000:], e[-2000:])
        assert "Param:" not in o, "rank %d: the script found weights outside its tolerances:\n%s" % (r, o[-4000:])
    assert "[%d] Stepping" % (STEPS - 1) in outs[0][1]


SCRIPT = r"""
import runpy, sys, os
sys.path[:0] = [%(pkg)r, %(tests)r]
import apex
assert os.path.realpath(apex.__file__).startswith(os.path.realpath(%(pkg)r)), apex.__file__
sys.argv = [%(name)r + ".py"] + (["--local_rank", os.environ["LOCAL_RANK"]] if "LOCAL_RANK" in os.environ else []) + %(extra)r
runpy.run_module(%(name)r, run_name="__main__")
"""


def _run_script(name, world, cwd, extra=(), env_extra=None):
    """Runs oracle/_ref/pyref_tests.zip/distributed/<name>.pyc as __main__ on `world` ranks (0 = plain single process)."""
    if not os.path.exists(TESTS_ZIP):
        pytest.skip("oracle/_ref/pyref_tests.zip not staged (oracle/build_ref.sh)")
    import zipfile
    if "distributed/%s.pyc" % name not in zipfile.ZipFile(TESTS_ZIP).namelist():
        pytest.skip("oracle/_ref/pyref_tests.zip predates the distributed scripts (re-run oracle/build_ref.sh)")
    if torch.cuda.device_count() < max(world, 1):
        pytest.skip("needs %d GPUs" % world)
    body = SCRIPT % dict(pkg=os.path.join(ROOT, "rocm-apex_b200"), tests=os.path.join(TESTS_ZIP, "distributed"), name=name,
                         extra=list(extra))
    port = _free_port()
    procs = []
    for r in range(max(world, 1)):
        env = dict(os.environ)
        env.pop("PYTHONPATH", None)
        for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        if world:
            env.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world))
        env.update(env_extra or {})
        procs.append(subprocess.Popen([sys.executable, "-c", body], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                      text=True, cwd=str(cwd)))
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, o, e))
    for r, (rc, o, e) in enumerate(outs):
        assert rc == 0, "%s rank %d left with %s\n%s\n%s" % (name, r, rc, o[-3000:], e[-4000:])
    return outs


def test_reference_ddp_race_condition_script_two_ranks(tmp_path):
    """tests/distributed/DDP/ddp_race_condition_test.py: apex.parallel.DistributedDataParallel with message_size=1,
    allreduce_trigger_params and three all-reduce streams; ten iterations whose gradient sums it checks exactly."""
    outs = _run_script("ddp_race_condition_test", 2, tmp_path)
    for r, (rc, o, e) in enumerate(outs):
        assert "passed =  True" in o, "rank %d:\n%s\n%s" % (r, o[-3000:], e[-2000:])
        assert o.count("expected sum") == 20


def test_reference_amp_master_params_scripts_two_ranks(tmp_path):
    """tests/distributed/amp_master_params: 500 amp-O2 SGD iterations under apex DDP on two ranks, then the reference's compare.py:
    model and master parameters identical across ranks, model == master within fp16."""
    _run_script("amp_master_params", 2, tmp_path)
    for f in ("rank0model.pth", "rank1model.pth", "rank0master.pth", "rank1master.pth"):
        assert (tmp_path / f).exists()
    outs = _run_script("compare", 0, tmp_path, env_extra={"TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD": "1"})
    assert "OK:  Model and master params match across ranks." in outs[0][1]


def test_reference_dist_adam_script_one_rank():
    _run(1, dim=4, layers=2)


def test_reference_dist_adam_script_one_rank_wider_model_with_bias():
    _run(1, dim=64, layers=3, extra=["--bias"])


def test_reference_dist_adam_script_two_ranks():
    _run(2, dim=32, layers=2, extra=["--bias"])
