mkdir -p gpurun_out
python tools/bench_configs.py --out gpurun_out/configs_final.md > gpurun_out/configs_final.jsonl 2> gpurun_out/configs_final.err; tail -3 gpurun_out/configs_final.err; wc -l gpurun_out/configs_final.md gpurun_out/configs_final.jsonl
python tools/sweep_variants.py > gpurun_out/variants_final.json 2>&1; tail -1 gpurun_out/variants_final.json
