export B200MT_BENCH_SKIP_CPU=1
python bench.py --steps 5 --warmup 3 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 966 -c 60 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
python bench.py --steps 5 --warmup 3 > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"AdamOp|ScaleOp" -s 6 -c 4 -o gpurun_out/adam_scale_r01 python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_full.log | cut -c1-200
