"""CPU: the reference arm of bench.py (`--impl reference`) needs no GPU: it must print ONE JSON line with the keys the bench
contract names, time the reference's own python-only unscale when oracle/_ref/pyref.zip is staged (cpu_baseline.kind
"reference"; the oracle port otherwise), and, under a multi-rank launch, print from rank 0 only."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "PYTHONPATH")}
    env.update(env_extra)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", *args], capture_output=True,
                       text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    return [x for x in r.stdout.splitlines() if x.startswith("{")]


def test_reference_arm_prints_one_contract_line():
    lines = _run({}, "--steps", "2", "--warmup", "1", "--max-seconds", "10")
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "fused_adam_step_params_per_s" and d["unit"] == "params/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["config"]["params"] == 25557032 and d["config"]["tensors"] == 161
    c = d["cpu_baseline"]
    staged = os.path.exists(os.path.join(ROOT, "oracle", "_ref", "pyref.zip"))
    assert c["kind"] == ("reference" if staged else "port") and c["cores"] >= 1 and c["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "params/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_under_a_two_rank_launch_prints_from_rank_0_only():
    assert _run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, "--gpus", "2", "--steps", "1", "--warmup", "0") == []
    lines = _run({"RANK": "0", "LOCAL_RANK": "0", "WORLD_SIZE": "2", "OMP_NUM_THREADS": "1"}, "--gpus", "2", "--steps", "1",
                 "--warmup", "0", "--max-seconds", "5")
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["config"]["params"] == 8030261248
    assert "sample" in d["cpu_baseline"] and d["cpu_baseline"]["cores"] >= 1


def test_own_arm_refuses_to_run_without_a_gpu():
    """The product path has no CPU fallback: on a box without a CUDA device `python bench.py` must say so and fail."""
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1"], capture_output=True,
                       text=True, timeout=300, cwd=ROOT)
    assert r.returncode != 0 and "no CPU fallback" in (r.stdout + r.stderr)
    assert not [x for x in r.stdout.splitlines() if x.startswith("{")]
