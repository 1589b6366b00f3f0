timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo rc=$? >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
export B200MT_BENCH_SKIP_CPU=1
rm -f gpurun_out/sweep_vec.log
for L in libb200mt.so libb200mt_v4u4.so libb200mt_v8u1.so libb200mt_v8u2.so; do
  for rep in 1 2; do
  B200MT_LIB=$PWD/rocm-apex_b200/$L python bench.py --steps 300 --warmup 10 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.readline())
print('$L step_us=%.1f adam_us=%.1f adam_frac=%.3f value=%.3e'%(d['ms_per_step']*1e3, d['roofline']['avg_launch_ms']*1e3, d['roofline']['frac'], d['value']))" >> gpurun_out/sweep_vec.log
  done
done
cat gpurun_out/sweep_vec.log
