python tools/kbench.py --keyframes 64 --device-scenes --plans 2
for v in l2 l4 m2 m4 m8 l2m4; do echo "== $v"; TDK_B200_LIB=tadataka_b200/libtdk_v_$v.so python tools/kbench.py --keyframes 64 --device-scenes --plans 2; done
