mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3; }
run 29813 > gpurun_out/sh_plain_fused_$N.log 2>&1; tail -1 gpurun_out/sh_plain_fused_$N.log | python -c "
import sys,json; d=json.loads(sys.stdin.readline()); print(d['n_gpus'], d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['gpu_launches'], d['clocks'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29815 bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-300
