python -m pytest tests -m gpu -q 2>&1 | tail -3
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -1 gpurun_out/final_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_ref.json 2> /dev/null
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'update_depth|key_prep|pair_setup' -c 36 --csv --log-file gpurun_out/final_launches.csv python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none --cache-control none -k regex:'update_depth|key_prep' -s 10 -c 5 -o gpurun_out/final_full python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > /dev/null 2>&1
ncu -i gpurun_out/final_full.ncu-rep --page raw --csv > gpurun_out/final_full_raw.csv 2>/dev/null
python tools/kbench.py --keyframes 64 --device-scenes --plans 2
ls -la gpurun_out/final_*
