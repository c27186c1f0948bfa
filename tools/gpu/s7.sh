tools/micro/l2_peak
python tools/kbench.py --keyframes 64 --device-scenes --plans 2
python tools/kbench.py --keyframes 64 --distinct 8 --plans 2,0,1
python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --c4
python -m pytest tests -m gpu -q -x 2>&1 | tail -4
