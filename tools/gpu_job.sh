mkdir -p gpurun_out
python -m pytest tests/test_reversible_gpu.py tests/test_dist_lamb.py tests/test_parity_more_gpu.py -m gpu -q > gpurun_out/pytest_rev.log 2>&1; tail -15 gpurun_out/pytest_rev.log
