mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
print("cpus allowed", len(os.sched_getaffinity(0)), "of", os.cpu_count())
for i in range(torch.cuda.device_count()):
    c = bench.gpu_numa_cpus(i)
    print("gpu", i, "near cpus", None if c is None else (min(c), max(c), len(c)))
PY
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3; }
run 29813 > gpurun_out/sh_plain_fused_$N.log 2>&1; tail -1 gpurun_out/sh_plain_fused_$N.log | python -c "
import sys,json; d=json.loads(sys.stdin.readline()); print(d['n_gpus'], d['ms_per_step'], d['e2e'])"
