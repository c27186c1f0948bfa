python -m pytest tests/test_gpu_update_depth.py -m gpu -q -k "host_pointer" 2>&1 | tail -3
python bench.py --steps 30 --no-c4 --c5-pairs 256 --no-cpu-baseline > gpurun_out/r2_bench_quick.json 2> gpurun_out/r2_bench_quick.err
tail -c 1500 gpurun_out/r2_bench_quick.json; tail -3 gpurun_out/r2_bench_quick.err
python tools/kbench.py --keyframes 64 --device-scenes --plans 2
