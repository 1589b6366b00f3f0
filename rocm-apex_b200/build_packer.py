"""Builds _b200mt_pack (csrc/packer.cpp): the torch-aware native walker of tensor lists.  No kernels, no CUDA: plain g++
against torch's headers, in-tree next to libb200mt.so.  Optional: _b200mt.py falls back to its Python walker without it.

  python rocm-apex_b200/build_packer.py
"""
import os
import subprocess
import sys
import sysconfig


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    import torch
    tdir = os.path.dirname(torch.__file__)
    suffix = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    out = os.path.join(here, "_b200mt_pack" + suffix)
    src = os.path.join(here, "csrc", "packer.cpp")
    if os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(src) and not os.environ.get("FORCE"):
        print("[build_packer] %s up to date" % os.path.basename(out))
        return 0
    abi = int(getattr(torch._C, "_GLIBCXX_USE_CXX11_ABI", True))
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-D_GLIBCXX_USE_CXX11_ABI=%d" % abi, src, "-o", out,
           "-I" + os.path.join(tdir, "include"), "-I" + os.path.join(tdir, "include", "torch", "csrc", "api", "include"),
           "-I" + sysconfig.get_paths()["include"], "-L" + os.path.join(tdir, "lib"), "-ltorch", "-ltorch_cpu", "-ltorch_python",
           "-lc10", "-Wl,-rpath," + os.path.join(tdir, "lib")]
    print("[build_packer] " + " ".join(cmd))
    subprocess.check_call(cmd)
    return 0


if __name__ == "__main__":
    sys.exit(main())
