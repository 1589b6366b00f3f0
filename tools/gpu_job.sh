mkdir -p gpurun_out
python tools/sweep_math.py > gpurun_out/math_approx.json 2> gpurun_out/math_approx.err; tail -1 gpurun_out/math_approx.json; tail -3 gpurun_out/math_approx.err
B200MT_LIB=$PWD/rocm-apex_b200/libb200mt_ieee.so python tools/sweep_math.py > gpurun_out/math_ieee.json 2> gpurun_out/math_ieee.err; tail -1 gpurun_out/math_ieee.json; tail -3 gpurun_out/math_ieee.err
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_math.log 2>&1; tail -1 gpurun_out/bench_math.log | python -c "
import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['mixed_precision_adam'], d['e2e']['value'])"
