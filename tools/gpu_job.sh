mkdir -p gpurun_out
python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
for i in range(torch.cuda.device_count()):
    c = bench.gpu_numa_cpus(i)
    print("gpu", i, "numa cpus", None if c is None else (min(c), max(c), len(c)), "affinity", len(os.sched_getaffinity(0)))
PY
for i in 1 2; do B200MT_BENCH_SKIP_CPU=1 python bench.py --steps 20 --warmup 3 > gpurun_out/b_numa_$i.log 2>&1; tail -1 gpurun_out/b_numa_$i.log | python -c "
import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['e2e']['value'])"; done
python tools/e2e_probe.py
