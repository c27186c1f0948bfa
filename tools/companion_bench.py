"""Times the companion kernels on P VGA frames and reports them against the HBM roofline
(algorithmic bytes per pixel from SURVEY 8d / BASELINE.md 4).
    python tools/companion_bench.py [--frames 64] [--iters 20]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tadataka_b200 import device, scene  # noqa: E402

def run(frames=64, iters=20, verbose=False):
    """-> {name: {"us_per_frame", "bytes_per_px", "gbs", "frac_of_hbm_peak"}} on `frames` VGA frames."""
    P, H, W = frames, 480, 640
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    rng = np.random.default_rng(0)
    base = scene.make_pair(1000)
    gt = base["gt_depth"]
    depth = torch.from_numpy(np.stack([gt * (1 + 0.05 * rng.standard_normal((H, W))) for _ in range(P)])).cuda()
    var = torch.from_numpy(rng.uniform(1e-6, 1e-4, (P, H, W))).cuda()
    var2 = torch.from_numpy(rng.uniform(1e-6, 1e-4, (P, H, W))).cuda()
    depth2 = depth * 1.01
    flag = torch.from_numpy(np.where(rng.random((P, H, W)) < 0.15, 0, -6).astype(np.int64)).cuda()
    age = torch.from_numpy(rng.integers(0, 6, (P, H, W)).astype(np.uint64)).cuda()
    cams = torch.from_numpy(np.tile(scene.CAM_VGA, (P, 1))).cuda()
    T10 = torch.from_numpy(np.stack([scene.pose(scene.rotation(0.01, -0.02, 0.005), (3.0, -1.0, 10.0 * (i % 5)))
                                     for i in range(P)])).cuda()

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    rows = [
        ("fusion", 48, lambda: device.fusion(depth, depth2, var, var2)),
        ("fusion_gated_depth_domain", 56, lambda: device.fusion(depth, depth2, var, var2, flag=flag, depth_domain=True)),
        ("regularize", 32, lambda: device.regularize(depth, var, flag)),
        ("increment_age", 24, lambda: device.increment_age(age, cams, cams, T10, depth)),
        ("propagate", 32, lambda: device.propagate(T10, cams, cams, depth, var, 200.0, 100.0, 1e-6)),
        ("propagate_with_age", 48, lambda: device.propagate_with_age(T10, cams, cams, depth, var, age, 200.0, 100.0, 1e-6)),
    ]
    out = {}
    if verbose:
        print(f"{P} x {W}x{H} frames, HBM peak {peak} GB/s")
    for name, bpp, fn in rows:
        ms = timeit(fn)
        gbs = bpp * P * H * W / ms / 1e6
        out[name] = {"us_per_frame": ms * 1e3 / P, "bytes_per_px": bpp, "gbs": gbs, "frac_of_hbm_peak": gbs / peak}
        if verbose:
            print(f"{name:32s} ({bpp} B/px) {ms * 1e3 / P:8.2f} us/frame  {P * H * W / ms / 1e6:7.2f} Gpx/s  {gbs:7.1f} GB/s  {100 * gbs / peak:5.1f}% of HBM peak")
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    run(args.frames, args.iters, verbose=True)
