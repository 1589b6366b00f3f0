"""CPU: the C-ABI library loads and exports every symbol include/b200mt.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200mt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200mt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    for must in ("b200mt_scale", "b200mt_axpby", "b200mt_l2norm", "b200mt_adam", "b200mt_sgd", "b200mt_lamb",
                 "b200mt_lamb_mp", "b200mt_dist_adam", "b200mt_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    path = os.path.join(ROOT, "rocm-apex_b200", "libb200mt.so")
    assert os.path.exists(path), "libb200mt.so not built: run __graft_entry__.build()"
    lib = ctypes.CDLL(path)
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing
    lib.b200mt_version.restype = ctypes.c_int
    assert lib.b200mt_version() >= 100


def test_python_binding_covers_header():
    import _b200mt
    assert sorted(_b200mt.EXPORTS) == [n for n in _declared()]


def test_bad_arguments_are_rejected_without_a_gpu():
    import _b200mt
    L = _b200mt.Lists()
    L.ntensors = 0
    L.depth = 2
    rc = _b200mt.lib.b200mt_scale(ctypes.byref(L), 65536, None, 1.0, None)
    assert rc != 0
    assert b"size() is not > 0" in _b200mt.lib.b200mt_last_error()
    rc = _b200mt.lib.b200mt_scale(ctypes.byref(L), 0, None, 1.0, None)
    assert rc != 0 and b"chunk_size" in _b200mt.lib.b200mt_last_error()


def test_amp_c_module_has_reference_entry_points():
    import amp_C
    for name in ("multi_tensor_scale", "multi_tensor_axpby", "multi_tensor_l2norm", "multi_tensor_l2norm_mp",
                 "multi_tensor_l2norm_scale", "multi_tensor_adam", "multi_tensor_sgd", "multi_tensor_lamb",
                 "multi_tensor_lamb_mp", "multi_tensor_lamb_stage1_cuda", "multi_tensor_lamb_stage2_cuda",
                 "multi_tensor_adagrad", "multi_tensor_novograd"):
        assert callable(getattr(amp_C, name))


def test_plain_c_host_compiles_links_and_gets_errors(tmp_path):
    """tests/cabi/host.c: a host in plain C99 (-Wall -Wextra -Werror -pedantic) includes the header, links the library
    and exercises argument checking -- no GPU involved."""
    import subprocess
    libdir = os.path.join(ROOT, "rocm-apex_b200")
    exe = str(tmp_path / "cabi_host")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cabi", "host.c"), "-o", exe, "-L", libdir, "-lb200mt",
                           "-Wl,-rpath," + libdir])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout + r.stderr
