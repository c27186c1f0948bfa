python -m pytest tests/test_gpu_companions.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -3
