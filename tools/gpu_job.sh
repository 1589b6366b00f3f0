mkdir -p gpurun_out
python -m pytest tests/test_optimizer_cache_gpu.py tests/test_optimizers_gpu.py tests/test_amp_more_gpu.py -m gpu -x -q > gpurun_out/pytest_opt.log 2>&1; tail -12 gpurun_out/pytest_opt.log
python tools/bench_optimizer_api.py 2>&1 | tail -1
