timeout 600 python -m pytest tests/test_optimizers_gpu.py -m gpu -q > gpurun_out/pytest_opt.log 2>&1; echo rc=$? >> gpurun_out/pytest_opt.log
export B200MT_BENCH_SKIP_CPU=1
for T in 2048 4096 8192 16384 32768 65536; do
  for C in 0 592 1184; do
    B200MT_TILE_ELEMS=$T B200MT_GRID_CAP=$C python bench.py --steps 200 --warmup 10 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.readline())
print('tile=$T cap=$C step_us=%.1f adam_us=%.1f adam_frac=%.3f value=%.3e'%(d['ms_per_step']*1e3, d['roofline']['avg_launch_ms']*1e3, d['roofline']['frac'], d['value']))" >> gpurun_out/sweep_tiles.log
  done
done
python bench.py --steps 5 --warmup 3 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 966 -c 60 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
tail -30 gpurun_out/pytest_opt.log; cat gpurun_out/sweep_tiles.log
