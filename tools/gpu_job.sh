mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_full.log 2>&1; tail -1 gpurun_out/bench_full.log | python -c "
import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['cuda_graph_replay'], d['cpu_baseline']['value'], d['gpu_launches'], d['table_uploads'], d['table_hits'])"; tail -3 gpurun_out/bench_full.log | head -2 | cut -c1-300
