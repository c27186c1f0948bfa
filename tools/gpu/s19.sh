python tools/kbench.py --keyframes 64 --device-scenes --plans 2
for v in prep4096 prep8192 prep16384; do echo "== $v"; TDK_B200_LIB=tadataka_b200/libtdk_v_$v.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2; done
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:'key_prep' -c 8 --csv --log-file gpurun_out/prep_a.csv python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > /dev/null 2>&1
TDK_B200_LIB=tadataka_b200/libtdk_v_prep8192.so ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:'key_prep' -c 8 --csv --log-file gpurun_out/prep_b.csv python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > /dev/null 2>&1
grep -h "key_prep" gpurun_out/prep_a.csv | tail -2 | cut -d, -f15-; grep -h "key_prep" gpurun_out/prep_b.csv | tail -2 | cut -d, -f15-
