"""Times one Gauss-Newton iteration of the direct pose-change estimator at 640x480 (the two
per-pixel passes of csrc/dvo.cu), against the HBM roofline and the numpy oracle on the host.
    python tools/dvo_bench.py [--iters 50]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import dvo as odvo  # noqa: E402
from tadataka_b200 import dvo, scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=50)
args = ap.parse_args()
H, W = 480, 640
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0
cam = scene.CAM_VGA
world = scene.World(scene.SET_S, 9100)
T_w1 = scene.pose(scene.rotation(0.004, -0.006, 0.003), (1.5, -0.6, 0.9))
I0, D0 = world.render(scene.pose(), cam, H, W)
I1, _ = world.render(T_w1, cam, H, W)
T10 = np.linalg.inv(T_w1)
level = dvo._Level(cam, cam, I0, D0, I1)


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / args.iters


rows = [("normal equations (40 B/px: I0, I1, D0 + 2 gradient gathers)", 40, lambda: level.normal_equations(T10, None)),
        ("normal equations, huber", 40, lambda: level.normal_equations(T10, "huber")),
        ("photometric error (24 B/px: I0, D0 + I1 gather)", 24, lambda: level.photometric_error(T10))]
print(f"{W}x{H}, HBM peak {PEAK} GB/s; times include the 240-byte result read-back (one sync per call)")
for name, bpp, fn in rows:
    ms = timeit(fn)
    print(f"{name:66s} {ms * 1e3:8.1f} us  {bpp * H * W / ms / 1e6:7.1f} GB/s  {100 * bpp * H * W / ms / 1e6 / PEAK:5.1f}% of HBM peak")
t0 = time.perf_counter()
T = dvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), None, max_iter=20)
torch.cuda.synchronize()
t_gpu = time.perf_counter() - t0
trace = []
t0 = time.perf_counter()
oT = odvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), None, max_iter=20, trace=trace)
t_cpu = time.perf_counter() - t0
print(f"full single-level estimate ({len(trace)} iterations): GPU {t_gpu * 1e3:.1f} ms, numpy oracle {t_cpu * 1e3:.1f} ms, "
      f"max |dT| {np.max(np.abs(T - oT)):.2e}")

# the loop alone, images already on the device: per-iteration host round trips against tdk_dvo_iterate_dev
for name, T_w, in (("small motion", T_w1), ("larger motion", scene.pose(scene.rotation(0.012, -0.018, 0.009), (6.0, -2.5, 3.5)))):
    J1, _ = world.render(T_w, cam, H, W)
    lv = dvo._Level(cam, cam, I0, D0, J1)
    for w in (None, "huber"):
        tr = []
        dvo._iterate(lv, np.eye(4), w, 20, tr)
        dvo._iterate(lv, np.eye(4), w, 20)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            Th = dvo._iterate(lv, np.eye(4), w, 20, [])
        t_host = (time.perf_counter() - t0) / 10
        t0 = time.perf_counter()
        for _ in range(10):
            Td = dvo._iterate(lv, np.eye(4), w, 20)
        t_dev = (time.perf_counter() - t0) / 10
        print(f"{name}, weights {w}: {len(tr)} iterations; host loop {t_host * 1e3:.2f} ms, device loop {t_dev * 1e3:.2f} ms, "
              f"max |dT| {np.max(np.abs(Th - Td)):.1e}")

# the coarse-to-fine estimator on one VGA pair, 5 levels: pyramid made on the device against the numpy one
from rust_bindings.camera import CameraParameters  # noqa: E402
cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
est = dvo.PoseChangeEstimator(cp, cp, n_coarse_to_fine=5, max_iter=20)
J1, _ = world.render(scene.pose(scene.rotation(0.012, -0.018, 0.009), (6.0, -2.5, 3.5)), cam, H, W)
gI0, gD0, gJ1 = (torch.from_numpy(a).cuda() for a in (I0, D0, J1))
for name, args in (("numpy images in", (I0, D0, J1)), ("CUDA tensors in", (gI0, gD0, gJ1))):
    est(*args)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        T = est(*args)
    t = (time.perf_counter() - t0) / 5
    print(f"PoseChangeEstimator, 5 levels, {name}: {t * 1e3:.2f} ms per call")
t0 = time.perf_counter()
Tn = np.eye(4)
for level in reversed(range(5)):
    sc = 1 / pow(1.5, level)
    lv = dvo._Level(cam * sc, cam * sc, dvo.rescale(I0, sc), dvo.rescale(D0, sc), dvo.rescale(J1, sc))
    Tn = dvo._iterate(lv, Tn, None, 20)
print(f"the same with the pyramid made by numpy / scipy on the host: {(time.perf_counter() - t0) * 1e3:.1f} ms; max |dT| {np.max(np.abs(T - Tn)):.1e}")
