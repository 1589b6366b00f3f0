W=${W:-2}
for M in bert_large llama3_8b; do
MODEL=$M timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29861 tools/bench_dist_lamb.py 2>&1 | grep -E '^\{|Error' >> gpurun_out/dist_lamb_n$W.log
done
cat gpurun_out/dist_lamb_n$W.log
