timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 tools/pcie_probe.py > gpurun_out/r2_pcie8.json 2> gpurun_out/r2_pcie8.err
echo rc=$?; tail -c 3000 gpurun_out/r2_pcie8.json
timeout 300 python bench.py --gpus 8 --steps 30 --no-c4 > gpurun_out/r2_bench_8gpu_b.json 2> gpurun_out/r2_bench_8gpu_b.err
echo rc=$?
python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_8gpu_b.json"))
print(d["value"]/1e9, d["ms_per_step"], json.dumps(d["config"]["nccl_gather"])[-600:], d["e2e"]["value"]/1e9, d["config"]["c5"]["total_ms"], d["config"]["c5"]["compute_only_ms"])
PY
