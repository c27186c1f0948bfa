python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/r2_t2.log
python tools/kbench.py --keyframes 64 --distinct 8 --plans 2,0,1 > gpurun_out/r2_kb2.log 2>&1
for v in p1c8 p2c6 p2c5 p4c6 p4c4 L3 L5 L6 M3 M5 MB64 MB256 S32 S16; do
  echo "== $v" >> gpurun_out/r2_kb2.log
  TDK_B200_LIB=tadataka_b200/libtdk_v_$v.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2 >> gpurun_out/r2_kb2.log 2>&1
done
echo "== default (device scenes)" >> gpurun_out/r2_kb2.log
python tools/kbench.py --keyframes 64 --device-scenes --plans 2 >> gpurun_out/r2_kb2.log 2>&1
python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --c4 >> gpurun_out/r2_kb2.log 2>&1
echo "== gate" >> gpurun_out/r2_kb2.log
TDK_B200_LIB=tadataka_b200/libtdk_v_gate.so python tools/kbench.py --keyframes 64 --distinct 8 --plans 2 >> gpurun_out/r2_kb2.log 2>&1
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2_launch2.csv python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > gpurun_out/r2_ncu2.log 2>&1
cat gpurun_out/r2_t2.log gpurun_out/r2_kb2.log
