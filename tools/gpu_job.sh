mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3; }
B200MT_SHARDED_FUSED=0 run 29814 > gpurun_out/sh_plain_nccl_$N.log 2>&1; tail -1 gpurun_out/sh_plain_nccl_$N.log | cut -c1-300
run 29813 > gpurun_out/sh_plain_fused_$N.log 2>&1; tail -1 gpurun_out/sh_plain_fused_$N.log | cut -c1-300
