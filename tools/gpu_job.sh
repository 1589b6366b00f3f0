# Scratch job for `gpurun -- 'bash tools/gpu_job.sh'`: the single-GPU verification pass (tests, bench, smoke).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; tail -1 gpurun_out/bench_full.json | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
