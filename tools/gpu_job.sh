mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python tools/bench_configs.py --out gpurun_out/configs_final.md > gpurun_out/configs_final.jsonl 2> gpurun_out/configs_final.err; tail -2 gpurun_out/configs_final.err; wc -l gpurun_out/configs_final.md
python bench.py > gpurun_out/bench_full.log 2>&1; tail -1 gpurun_out/bench_full.log | cut -c1-200
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
