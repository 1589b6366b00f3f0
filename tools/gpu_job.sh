mkdir -p gpurun_out
python -m pytest tests/test_adam_tma_gpu.py -m gpu -x -q 2>&1 | tail -4
