mkdir -p gpurun_out
for t in 3072 3584 3840 4000 4096 4352 4608 5120 6144; do B200MT_TILE_ELEMS=$t B200MT_BENCH_SKIP_CPU=1 python bench.py --steps 100 --warmup 5 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.readline()); print($t, round(d['ms_per_step']*1e3,2), round(d['roofline']['avg_launch_ms']*1e3,2), round(d['roofline']['frac'],4), round(d['mixed_precision_adam']['ms_per_launch']*1e3,2))"; done
