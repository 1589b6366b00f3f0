timeout 900 python -m pytest tests/test_amp_more_gpu.py -m gpu -q > gpurun_out/pytest_amp_more.log 2>&1; echo rc=$? >> gpurun_out/pytest_amp_more.log
grep -E "^E  |passed|failed|Error|^tests.*(Error|assert)" gpurun_out/pytest_amp_more.log | head -30
