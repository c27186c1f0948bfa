for v in pipe pipe3 un2 un2c3; do echo "== $v"; TDK_B200_LIB=tadataka_b200/libtdk_v_$v.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2; done
python tools/kbench.py --keyframes 64 --device-scenes --plans 2
TDK_B200_LIB=tadataka_b200/libtdk_v_pipe.so python tools/kbench.py --keyframes 16 --distinct 4 --plans 2
