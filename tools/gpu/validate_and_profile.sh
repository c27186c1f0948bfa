# every GPU test, the default bench line, the reference arm, then the two ncu captures of one warm
# 64-keyframe step that profiles/<commit>_* are made from (run through tools/gpu/retry.sh)
python -m pytest tests -m gpu -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.build(); g.smoke()"
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -1 gpurun_out/final_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_ref.json 2> /dev/null
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'update_depth|key_prep|pair_setup' -c 36 --csv --log-file gpurun_out/final_launches.csv python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none --cache-control none -k regex:'update_depth|key_prep' -s 10 -c 5 -o gpurun_out/final_full python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > /dev/null 2>&1
ncu -i gpurun_out/final_full.ncu-rep --page raw --csv > gpurun_out/final_full_raw.csv 2>/dev/null
python tools/kbench.py --keyframes 64 --device-scenes --plans 2
python -c "
import json; d=json.load(open('gpurun_out/final_bench.json')); c=d['config']
print(d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['e2e']['value']/1e9, c['c3_vo_loop_ms_per_frame'], c['c4_720p_ms_per_pair'], c['c5']['total_ms'], c['c5']['checks'], d['cpu_baseline']['value']/1e6, d['cpu_baseline']['maps_identical_to_gpu'])"
