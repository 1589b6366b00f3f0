"""GPU: a host program in plain C (tests/cabi/host_gpu.c) drives the kernels through include/b200mt.h and the CUDA runtime
alone -- cudaMalloc'd buffers, a cudaStream_t, no torch, no C++ -- and checks unscale + overflow flag, global and per-tensor
norm and an AdamW step (through a prepared handle) against the same arithmetic done in C.  This is the drop-in boundary of
SURVEY.md §8b exercised the way a non-Python host would."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plain_c_host_runs_scale_norm_and_adam_on_the_gpu(tmp_path):
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    libdir = os.path.join(ROOT, "rocm-apex_b200")
    exe = str(tmp_path / "cabi_host_gpu")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-I", os.path.join(cuda, "include"), os.path.join(ROOT, "tests", "cabi", "host_gpu.c"), "-o", exe,
                           "-L", libdir, "-lb200mt", "-Wl,-rpath," + libdir, "-L", os.path.join(cuda, "lib64"), "-lcudart", "-lm"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-3000:] + r.stderr[-2000:]
