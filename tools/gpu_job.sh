timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo rc=$? >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
timeout 1200 python tools/bench_configs.py --out gpurun_out/configs.md > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo rc=$?; tail -2 gpurun_out/configs.err; cut -c1-230 gpurun_out/configs.md
