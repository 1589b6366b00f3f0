mkdir -p gpurun_out
python -m pytest tests/test_dist_adam_gpu.py -k "rank_groups or matches_unsharded" -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; tail -8 gpurun_out/pytest_gpu2.log
