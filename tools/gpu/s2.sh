python -m pytest tests/test_gpu_dvo.py -m gpu -q 2>&1 | tail -3 > gpurun_out/r2_t3.log; cat gpurun_out/r2_t3.log
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'update_depth|key_prep|pair_setup' -c 36 --csv --log-file gpurun_out/r2_launch2.csv python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > gpurun_out/r2_ncu2.log 2>&1
python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
tail -c 4000 gpurun_out/r2_bench.json; tail -5 gpurun_out/r2_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
tail -c 600 gpurun_out/r2_bench_ref.json
python tools/companion_bench.py > gpurun_out/r2_companions.log 2>&1
tail -12 gpurun_out/r2_companions.log
