python tools/kbench.py --keyframes 64 --device-scenes --plans 2
TDK_B200_LIB=tadataka_b200/libtdk_v_hack.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2
