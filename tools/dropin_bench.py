"""Latency of the numpy drop-in calls (rust_bindings.semi_dense.*: numpy in, numpy out, one call per
frame as in examples/semi_dense_vo.py:196-206) on one VGA keyframe against R reference frames.
    python tools/dropin_bench.py [--refs 8] [--iters 30]
Development tool; bench.py reports the same figures as config.dropin."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(refs=8, iters=30, verbose=False):
    import torch
    from rust_bindings import semi_dense as sd
    from rust_bindings.camera import CameraParameters
    from tadataka_b200 import scene
    it = scene.make_keyframe(2000, R=refs)
    cam = CameraParameters((it["key_cam"][0], it["key_cam"][1]), (it["key_cam"][2], it["key_cam"][3]))
    key = sd.Frame(cam, it["key_img"], it["T_wk"])
    frames = [sd.Frame(cam, it["ref_imgs"][r], it["T_wr"][r]) for r in range(refs)]
    params = sd.Params(*[float(x) for x in it["params"]])
    age, pd, pv = it["age"], it["prior_depth"], it["prior_variance"]
    T10 = np.linalg.inv(it["T_wk"]) @ it["T_wr"][-1]

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        torch.cuda.synchronize()
        return 1e3 * (time.perf_counter() - t0) / iters

    out = {
        "update_depth_ms": timeit(lambda: sd.update_depth(key, frames, age, pd, pv, params)),
        "propagate_ms": timeit(lambda: sd.propagate(T10, cam, cam, pd, pv, 200.0, 100.0, 1e-6)),
        "increment_age_ms": timeit(lambda: sd.increment_age(age, cam, cam, T10, pd)),
        "refs": refs, "shape": list(age.shape),
    }
    if verbose:
        print(out)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--refs", type=int, default=8)
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    run(a.refs, a.iters, verbose=True)
