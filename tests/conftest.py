import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")
    # A fresh checkout has no built artefacts (*.so is git-ignored): build the CUDA library (nvcc
    # cross-compiles sm_100a without a GPU) and the oracle before any test imports them.
    from tadataka_b200 import build as tdk_build
    if tdk_build.needs_build():
        tdk_build.build()
    import oracle  # noqa: F401  (compiles oracle/libtdk_oracle.so on first import)
