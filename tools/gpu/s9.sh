timeout 300 python bench.py --gpus 2 --steps 20 --no-c4 --c5-pairs 512 --e2e-steps 2 > gpurun_out/r2_g2_f.json 2> gpurun_out/r2_g2_f.err
echo rc=$?; tail -2 gpurun_out/r2_g2_f.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2_g2_f.json"))
print(d["ms_per_step"], json.dumps(d["config"]["nccl_gather"])[:900])
PY
