python -m pytest tests/test_gpu_companions.py tests/test_gpu_vo_loop.py -m gpu -q 2>&1 | tail -3
python tools/companion_bench.py
python bench.py > gpurun_out/r2_bench_b.json 2> gpurun_out/r2_bench_b.err; tail -2 gpurun_out/r2_bench_b.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_b.json"))
c=d["config"]
print(d["value"]/1e9, d["ms_per_keyframe"], d["roofline"]["frac"], d["e2e"]["value"]/1e9, c["c3_vo_loop_ms_per_frame"], c["c4_720p_ms_per_pair"], c["single_keyframe_launch_ms"], d["cpu_baseline"]["value"]/1e6, d["cpu_baseline"]["maps_identical_to_gpu"])
print({k:round(v["us_per_frame"],2) for k,v in c["companions"].items()})
PY
