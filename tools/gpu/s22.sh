timeout 300 python bench.py --gpus 2 --steps 10 --no-c4 --c5-pairs 512 --e2e-steps 1 > gpurun_out/r2_g2_g.json 2> gpurun_out/r2_g2_g.err; echo rc=$?
python -c "
import json; d=json.load(open('gpurun_out/r2_g2_g.json')); print(d['config']['c5']['checks'], d['config']['c5']['total_ms'])"
