timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo rc=$? >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_parity_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "misaligned or many_tensors or l2norm_reference_kat or lamb_mp_matches or noalloc_norm or adam_mixed or sgd_noop" > gpurun_out/memcheck.log 2>&1; echo memcheck_rc=$? >> gpurun_out/memcheck.log
tail -6 gpurun_out/memcheck.log
