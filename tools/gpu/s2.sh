# bench line (default), reference arm (short), companions
python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
tail -c 3500 gpurun_out/r2_bench.json; tail -5 gpurun_out/r2_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
tail -c 600 gpurun_out/r2_bench_ref.json
python tools/companion_bench.py > gpurun_out/r2_companions.log 2>&1
tail -20 gpurun_out/r2_companions.log
