python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_t1.log
python tools/kbench.py --keyframes 64 --distinct 8 --plans 2,0,1 > gpurun_out/r2_kb1.log 2>&1
python tools/kbench.py --keyframes 64 --distinct 8 --plans 2 --c4 --no-parity >> gpurun_out/r2_kb1.log 2>&1
TDK_B200_LIB=tadataka_b200/libtdk_b200_gate.so python tools/kbench.py --keyframes 64 --distinct 8 --plans 2 > gpurun_out/r2_kb1_gate.log 2>&1
tools/micro/l2_peak > gpurun_out/r2_l2peak.json 2>&1
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2_launch1.csv python tools/kbench.py --keyframes 64 --distinct 8 --plans 2 --no-parity --iters 2 > gpurun_out/r2_ncu1.log 2>&1
cat gpurun_out/r2_t1.log gpurun_out/r2_kb1.log gpurun_out/r2_kb1_gate.log gpurun_out/r2_l2peak.json
