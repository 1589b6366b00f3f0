timeout 900 python -m pytest tests/test_parity_more_gpu.py -m gpu -q > gpurun_out/pytest_more.log 2>&1; echo rc=$? >> gpurun_out/pytest_more.log
grep -E "^E  |passed|failed" gpurun_out/pytest_more.log | head -30
