timeout 900 python -m pytest tests/test_dist_lamb.py tests/test_properties_gpu.py tests/test_dist_adam_gpu.py -m gpu -q > gpurun_out/pytest_lamb.log 2>&1; echo rc=$? >> gpurun_out/pytest_lamb.log
grep -E "^E  |passed|failed|Error" gpurun_out/pytest_lamb.log | head -30
