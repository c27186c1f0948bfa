tools/micro/l2_peak
python -m pytest tests/test_gpu_vo_loop.py -m gpu -q 2>&1 | tail -4
python tools/vo_bench.py --plans 0 2>&1 | tail -3
python bench.py --steps 20 --c5-pairs 0 --no-cpu-baseline --c4-pairs 0 > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2_bench_c3.err; tail -2 gpurun_out/r2_bench_c3.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_c3.json"))
print("c3", d["config"]["c3_vo_loop_ms_per_frame"], d["config"]["c3_vo_loop_f64_frames_ms_per_frame"])
PY
