python tools/kbench.py --keyframes 64 --device-scenes --plans 2
TDK_B200_LIB=tadataka_b200/libtdk_v_tma.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2
TDK_B200_LIB=tadataka_b200/libtdk_v_tma6.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2
TDK_B200_LIB=tadataka_b200/libtdk_v_tma.so timeout 300 python tools/kbench.py --keyframes 16 --distinct 4 --plans 2,0
TDK_B200_LIB=tadataka_b200/libtdk_v_tma.so timeout 600 python -m pytest tests/test_gpu_update_depth.py -m gpu -q -x 2>&1 | tail -3
TDK_B200_LIB=tadataka_b200/libtdk_v_tma.so ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,lts__t_sectors.sum --clock-control none --cache-control none -k regex:sieve -s 3 -c 1 python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 2>&1 | grep -E "gpu__time|inst_executed|lts__"
