B200MT_PDL=1 timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_pdl.log 2>&1; echo rc=$? >> gpurun_out/pytest_gpu_pdl.log
tail -3 gpurun_out/pytest_gpu_pdl.log
export B200MT_BENCH_SKIP_CPU=1
rm -f gpurun_out/sweep_pdl.log
for P in 0 1 0 1; do
  B200MT_PDL=$P python bench.py --steps 300 --warmup 10 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.readline())
print('pdl=$P step_us=%.1f adam_us=%.1f adam_frac=%.3f value=%.3e'%(d['ms_per_step']*1e3, d['roofline']['avg_launch_ms']*1e3, d['roofline']['frac'], d['value']))" >> gpurun_out/sweep_pdl.log
done
cat gpurun_out/sweep_pdl.log
timeout 1200 python tools/bench_configs.py --out gpurun_out/configs.md > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo rc=$?; tail -3 gpurun_out/configs.err; cat gpurun_out/configs.md
