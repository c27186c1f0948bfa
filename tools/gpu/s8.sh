timeout 420 python bench.py --gpus 8 --steps 20 --no-c4 > gpurun_out/r2_bench_8gpu.json 2> gpurun_out/r2_bench_8gpu.err
echo "rc=$?"; tail -c 2600 gpurun_out/r2_bench_8gpu.json; tail -3 gpurun_out/r2_bench_8gpu.err
