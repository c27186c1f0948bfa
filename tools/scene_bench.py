"""Time the CUDA ray-caster (tdk_render_scene_dev) against the numpy renderer it mirrors.
    python tools/scene_bench.py [--views 72]
Development tool (not part of the bench contract)."""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tadataka_b200 import scene, scene_device  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--views", type=int, default=72)
args = ap.parse_args()
H, W = 480, 640
world = scene.World(scene.SET_S, 1)
rng = np.random.default_rng(0)
poses = np.stack([scene.pose(scene.rotation(*(rng.uniform(-0.01, 0.01, 3))), rng.uniform(-10, 10, 3))
                  for _ in range(args.views)])
cams = np.tile(scene.CAM_VGA, (args.views, 1))
scene_device.render(world, poses, cams, H, W)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    scene_device.render(world, poses, cams, H, W)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
t0 = time.time()
for i in range(3):
    world.render(poses[i], scene.CAM_VGA, H, W)
cpu_ms = (time.time() - t0) / 3 * 1e3
px = args.views * H * W
print(f"device: {ms:.3f} ms for {args.views} VGA views = {ms * 1e3 / args.views:.1f} us/view, "
      f"{px * 16 / ms / 1e6:.0f} GB/s written; numpy: {cpu_ms:.0f} ms/view")
