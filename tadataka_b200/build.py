"""Builds tadataka_b200/libtdk_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m tadataka_b200.build [--force] [--verbose]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libtdk_b200.so")
SOURCES = ["update_depth.cu", "companions.cu", "helpers.cu", "dvo.cu", "scene.cu", "mapper.cu", "capi.cu"]
HEADERS = [os.path.join(CSRC, "tdk_common.cuh"),
           os.path.join(os.path.dirname(HERE), "include", "tdk.h")]

# -fmad=false: parity with the reference needs every product and sum rounded separately.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false",
              "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, extra_flags=(), out=None):
    """extra_flags / out: development variants (e.g. -DTDK_MIN_CTAS=4 -> another .so name)."""
    if out is None and not force and not needs_build():
        return LIB
    env = dict(os.environ)
    # the image exports CC/CXX=/opt/gcc/..., nvcc wants the system host compiler
    cmd = [_nvcc(), "-ccbin", "/usr/bin/g++"] + NVCC_FLAGS
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += list(extra_flags)
    cmd += ["-o", out or LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libtdk_b200.so")
    return out or LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
