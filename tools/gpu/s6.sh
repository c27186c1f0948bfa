for mode in "d" "e"; do
  if [ $mode = d ]; then export NCCL_P2P_USE_CUDA_MEMCPY=1; fi
  if [ $mode = e ]; then export NCCL_P2P_USE_CUDA_MEMCPY=1; export CUDA_DEVICE_MAX_CONNECTIONS=32; fi
  python bench.py --gpus 2 --steps 20 --no-c4 --c5-pairs 512 --e2e-steps 1 > gpurun_out/r2_g2_$mode.json 2> gpurun_out/r2_g2_$mode.err
  tail -3 gpurun_out/r2_g2_$mode.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2_g2_$mode.json"))
print("$mode", "ms/step", round(d["ms_per_step"],3), "gather GB/s", round(d["config"]["nccl_gather"]["gbs_into_rank0"],1), "c5 total", round(d["config"]["c5"]["total_ms"],2), "compute", round(d["config"]["c5"]["compute_only_ms"],2), "c5 GB/s", round(d["config"]["c5"]["gather"]["gbs_into_rank0"],1))
PY
done
