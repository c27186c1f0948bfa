# two GPUs: sharded parity test, then the bench with the gather inside the clock
python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -8 > gpurun_out/r2_multi.log
cat gpurun_out/r2_multi.log
python bench.py --gpus 2 --steps 20 --no-c4 --c5-pairs 1024 > gpurun_out/r2_bench_2gpu.json 2> gpurun_out/r2_bench_2gpu.err
tail -c 3000 gpurun_out/r2_bench_2gpu.json; tail -5 gpurun_out/r2_bench_2gpu.err
