mkdir -p gpurun_out
rm -f gpurun_out/sweep_norm16_r02.txt
for v in n1 n2 n3 n4 n5; do B200MT_LIB=$PWD/rocm-apex_b200/libb200mt_$v.so python tools/sweep_light.py 2>>gpurun_out/sweep_n16.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(json.dumps({k:v for k,v in d.items() if k=='lib' or 'l2norm' in k}))" >> gpurun_out/sweep_norm16_r02.txt; done
python tools/sweep_light.py 2>>gpurun_out/sweep_n16.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(json.dumps({k:v for k,v in d.items() if k=='lib' or 'l2norm' in k}))" >> gpurun_out/sweep_norm16_r02.txt
cat gpurun_out/sweep_norm16_r02.txt
B200MT_LIB=$PWD/rocm-apex_b200/libb200mt_ieee.so python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -q > gpurun_out/pytest_full_ieee_r02.log 2>&1; tail -2 gpurun_out/pytest_full_ieee_r02.log
python -m pytest tests/test_parity_fullsize_gpu.py -m gpu -q > gpurun_out/pytest_full_r02.log 2>&1; tail -2 gpurun_out/pytest_full_r02.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
