#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- builds the *unmodified* reference kernels as a checker.
#
# Compiles the reference's own CUDA sources for the multi-tensor optimizer path, from where
# they lie under /root/reference (nothing is copied into this repo), for sm_100a, into
# oracle/_ref/ (git-ignored, but it travels to the GPU box with gpurun):
#   amp_C_ref.so               <- csrc/multi_tensor_*.cu + csrc/amp_C_frontend.cpp
#   distributed_adam_cuda_ref.so <- apex/contrib/csrc/optimizers/multi_tensor_distopt_adam*.{cpp,cu}
#   fused_adam_cuda_ref.so     <- apex/contrib/csrc/optimizers/fused_adam_cuda*.{cpp,cu}
#   distributed_lamb_cuda_ref.so <- apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb*.{cpp,cu}
# The reference's setup.py cannot be used (it shells out to /opt/rocm/bin/rocminfo at import,
# setup.py:15-17), so this is a direct nvcc/g++ recipe with the flags setup.py:210 uses
# (-O3 --use_fast_math -lineinfo + the VERSION_GE_* macros, setup.py:145-154) plus the four
# -D__CUDA_NO_HALF*/BFLOAT16* defines torch's CUDAExtension adds to every nvcc line (without them
# multi_tensor_distopt_lamb_kernel.cu:352 meets an ambiguous operator!= on at::Half).
#   pyref.zip, pyref_tests.zip <- the reference's Python package and its unit tests for the path, as bytecode archives
# Only tests/, __graft_entry__.smoke() and bench.py's comparison legs may load these.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
if [ ! -d "$REF/csrc" ]; then
  echo "[oracle/build_ref] $REF absent: keeping prebuilt oracle/_ref as is"; exit 0
fi
mkdir -p "$OBJ"
PY=${PYTHON:-python}
TORCH=$($PY -c 'import torch,os;print(os.path.dirname(torch.__file__))')
PYINC=$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')
INC="-I$REF/csrc -I$TORCH/include -I$TORCH/include/torch/csrc/api/include -I$PYINC -I/usr/local/cuda/include"
DEFS="-DVERSION_GE_1_1 -DVERSION_GE_1_3 -DVERSION_GE_1_5 -DTORCH_API_INCLUDE_EXTENSION_H"
NOHALF="-D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_BFLOAT16_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__"
NVFLAGS="$NOHALF -O3 -std=c++17 --use_fast_math -lineinfo --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -w"
LIBS="-L$TORCH/lib -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -ltorch_python -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$TORCH/lib"
JOBS=${JOBS:-8}

build_module () {   # name, then source files
  local name=$1; shift
  local objs=() pids=()
  if [ -f "$OUT/$name.so" ] && [ -z "${FORCE:-}" ]; then echo "[oracle/build_ref] $name.so up to date"; return; fi
  local running=0
  for src in "$@"; do
    local o="$OBJ/${name}_$(basename "${src%.*}").o"
    objs+=("$o")
    if [[ "$src" == *.cu ]]; then
      nvcc -c "$src" -o "$o" $INC $DEFS -DTORCH_EXTENSION_NAME=$name $NVFLAGS &
    else
      g++ -c "$src" -o "$o" $INC $DEFS -DTORCH_EXTENSION_NAME=$name -O2 -std=c++17 -fPIC -w &
    fi
    pids+=($!)
    running=$((running+1))
    if [ $running -ge $JOBS ]; then wait -n; running=$((running-1)); fi
  done
  for p in "${pids[@]}"; do wait $p; done
  g++ -shared "${objs[@]}" -o "$OUT/$name.so" $LIBS
  echo "[oracle/build_ref] built $OUT/$name.so"
}

build_module amp_C_ref \
  $REF/csrc/amp_C_frontend.cpp \
  $REF/csrc/multi_tensor_scale_kernel.cu $REF/csrc/multi_tensor_axpby_kernel.cu \
  $REF/csrc/multi_tensor_l2norm_kernel.cu $REF/csrc/multi_tensor_l2norm_kernel_mp.cu \
  $REF/csrc/multi_tensor_l2norm_scale_kernel.cu $REF/csrc/multi_tensor_adam.cu \
  $REF/csrc/multi_tensor_sgd_kernel.cu $REF/csrc/multi_tensor_lamb.cu \
  $REF/csrc/multi_tensor_lamb_mp.cu $REF/csrc/multi_tensor_lamb_stage_1.cu \
  $REF/csrc/multi_tensor_lamb_stage_2.cu $REF/csrc/multi_tensor_novograd.cu \
  $REF/csrc/multi_tensor_adagrad.cu
build_module distributed_adam_cuda_ref \
  $REF/apex/contrib/csrc/optimizers/multi_tensor_distopt_adam.cpp \
  $REF/apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu
build_module fused_adam_cuda_ref \
  $REF/apex/contrib/csrc/optimizers/fused_adam_cuda.cpp \
  $REF/apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu
build_module distributed_lamb_cuda_ref \
  $REF/apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp \
  $REF/apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu
rm -rf "$OBJ"

# The reference's own Python for this path, staged next to the compiled checkers AS BYTECODE ONLY (python -m compileall -b, the
# .py files are deleted again): like the .so files it is a build product of the reference, importable but not source, so that
# tests/test_reference_python_gpu.py and tests/test_reference_suite_gpu.py can run the reference's optimizer / amp classes and
# its own unit tests on the GPU box -- once over this repo's amp_C shim and once over amp_C_ref -- and compare.
# oracle/_ref/ is git-ignored: nothing of it enters the history.
STAGE=2   # bump when the set of staged files changes
if [ ! -f "$OUT/pyref.zip" ] || [ ! -f "$OUT/pyref_tests.zip" ] || [ "$(cat "$OUT/pyref.stage" 2>/dev/null)" != "$STAGE" ] || [ -n "${FORCE:-}" ]; then
  rm -rf "$OUT/pyref" "$OUT/pyref_tests" "$OUT/pyref.zip" "$OUT/pyref_tests.zip"
  mkdir -p "$OUT/pyref" "$OUT/pyref_tests"
  (cd "$REF" && find apex -name '*.py' | while read -r f; do mkdir -p "$OUT/pyref/$(dirname "$f")"; cp "$f" "$OUT/pyref/$f"; done)
  for d in run_amp run_optimizers; do
    mkdir -p "$OUT/pyref_tests/$d"
    cp "$REF"/tests/L0/$d/*.py "$OUT/pyref_tests/$d/"
  done
  # the reference's two-rank scripts for the DDP bucket path and amp master weights (tests/distributed/*)
  mkdir -p "$OUT/pyref_tests/distributed"
  cp "$REF"/tests/distributed/DDP/ddp_race_condition_test.py "$REF"/tests/distributed/amp_master_params/amp_master_params.py \
     "$REF"/tests/distributed/amp_master_params/compare.py "$OUT/pyref_tests/distributed/"
  $PY -W ignore -m compileall -q -b "$OUT/pyref" "$OUT/pyref_tests" > /dev/null 2>&1 || true   # files that do not compile under this Python are not on the path
  find "$OUT/pyref" "$OUT/pyref_tests" -name '*.py' -delete
  n=$(find "$OUT/pyref" "$OUT/pyref_tests" -name '*.pyc' | wc -l)
  # one importable archive each (zipimport reads sourceless .pyc): a single binary file travels with gpurun like the .so files
  (cd "$OUT/pyref" && $PY -m zipfile -c "$OUT/pyref.zip" apex)
  (cd "$OUT/pyref_tests" && $PY -m zipfile -c "$OUT/pyref_tests.zip" run_amp run_optimizers distributed)
  echo "$STAGE" > "$OUT/pyref.stage"
  rm -rf "$OUT/pyref" "$OUT/pyref_tests"
  echo "[oracle/build_ref] staged $n reference modules as bytecode in $OUT/pyref.zip, $OUT/pyref_tests.zip"
fi
