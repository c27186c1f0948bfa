timeout 400 python bench.py --gpus 4 --steps 30 > gpurun_out/r2_bench_4gpu.json 2> gpurun_out/r2_bench_4gpu.err
echo rc=$?; tail -2 gpurun_out/r2_bench_4gpu.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_4gpu.json"))
c=d["config"]
print(d["value"]/1e9, d["ms_per_step"], c["nccl_gather"]["depth_maps_only"]["value"]/1e9, c["nccl_gather"]["no_gather"]["value"]/1e9, d["e2e"]["value"]/1e9, c["c5"]["total_ms"], c["c5"]["compute_only_ms"], c["c4_720p_ms_per_pair"], c["c3_vo_loop_ms_per_frame"])
PY
