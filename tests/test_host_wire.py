"""Host half of tdk_update_depth_frames_host's byte wire (csrc/capi.cu): the worker pool narrows the
caller's u64 age maps to bytes and widens byte flags to i64.  Runs without a GPU."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

from tadataka_b200 import _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _round_trip(age, flag):
    age_back = np.full(age.shape, 12345, dtype=np.uint64)
    flag_back = np.full(flag.shape, 77, dtype=np.int64)
    _lib.check(_lib.selftest_host_wire(_p(age), _p(flag), age.size, _p(age_back), _p(flag_back)))
    return age_back, flag_back


def test_ages_saturate_and_flags_keep_their_sign():
    rng = np.random.default_rng(5)
    for n in (0, 1, 7, 76800, 76801, 3 * 76800 + 11, 640 * 480 * 2):
        age = rng.integers(0, 12, size=n).astype(np.uint64)
        flag = -rng.integers(0, 10, size=n).astype(np.int64)
        if n > 6:
            age[[1, 3, 5]] = [255, 256, 2 ** 40 + 1]       # all three must arrive as "more than any R < 255"
            age[-1] = 2 ** 64 - 1
            flag[[0, 2]] = [-128, 127]
        age_back, flag_back = _round_trip(age, flag)
        assert np.array_equal(age_back, np.minimum(age, 255))
        assert np.array_equal(flag_back, flag)


def test_pool_size_follows_the_environment():
    """TDK_HOST_THREADS = 1 and = 5 give the same answer (a fresh process each: the pool is made once)."""
    code = ("import numpy as np, ctypes as C\n"
            "from tadataka_b200 import _lib\n"
            "a = (np.arange(500000) % 300).astype(np.uint64); f = -(np.arange(500000) % 10).astype(np.int64)\n"
            "ab = np.empty_like(a); fb = np.empty_like(f)\n"
            "p = lambda x: x.ctypes.data_as(C.c_void_p)\n"
            "for _ in range(3): _lib.check(_lib.selftest_host_wire(p(a), p(f), a.size, p(ab), p(fb)))\n"
            "assert np.array_equal(ab, np.minimum(a, 255)) and np.array_equal(fb, f)\n"
            "print('ok')\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for threads in ("1", "5"):
        env = dict(os.environ, TDK_HOST_THREADS=threads, PYTHONPATH=root)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]
