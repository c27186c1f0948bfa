python -m pytest tests/test_gpu_companions.py tests/test_gpu_vo_loop.py -m gpu -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.build(); g.smoke()"
python tools/companion_bench.py | tail -3
