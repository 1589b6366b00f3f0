mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3; }
run 29813 > gpurun_out/sh_plain_fused_$N.log 2>&1; tail -1 gpurun_out/sh_plain_fused_$N.log | cut -c1-300
B200MT_SHARDED_E5M2=1 run 29811 > gpurun_out/sh_e5m2_fused_$N.log 2>&1; tail -1 gpurun_out/sh_e5m2_fused_$N.log | cut -c1-300
B200MT_SHARDED_E5M2=1 B200MT_SHARDED_FUSED=0 run 29812 > gpurun_out/sh_e5m2_nccl_$N.log 2>&1; tail -1 gpurun_out/sh_e5m2_nccl_$N.log | cut -c1-300
