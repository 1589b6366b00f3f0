mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python tools/sweep_variants.py 2>&1 | tail -1
python tools/bench_configs.py --skip-llama --out gpurun_out/configs_final.md > gpurun_out/configs_final.jsonl 2> gpurun_out/configs_final.err; grep "bert_large\|resnet50" gpurun_out/configs_final.md | cut -c1-200
python bench.py > gpurun_out/bench_full.log 2>&1; tail -1 gpurun_out/bench_full.log | cut -c1-250
