python -m pytest tests/test_gpu_companions.py tests/test_gpu_vo_loop.py -m gpu -q 2>&1 | tail -4
python tools/companion_bench.py | tail -4
