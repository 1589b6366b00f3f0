W=${W:-2}
timeout 600 python -m pytest tests/test_ddp.py -m gpu -q > gpurun_out/pytest_ddp.log 2>&1; echo rc=$? >> gpurun_out/pytest_ddp.log
grep -E "^E  |passed|failed|Error" gpurun_out/pytest_ddp.log | head -20
rm -f gpurun_out/ddp_n$W.log
for M in resnet50 bert_large; do
MODEL=$M timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29871 tools/bench_ddp.py 2>&1 | grep -E '^\{|Error' >> gpurun_out/ddp_n$W.log
done
cat gpurun_out/ddp_n$W.log
