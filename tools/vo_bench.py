"""C3: the semi_dense_vo keyframe loop (increment_age -> propagate -> update_depth + fusion ->
regularize) over a synthetic 640x480 sequence with ground-truth poses, device resident
(tadataka_b200.mapper.SemiDenseMapper).  Prints ms per frame for each kernel plan, and the share of
update_depth in it.
    python tools/vo_bench.py [--frames 30] [--plans 2,0]
Development tool (not part of the bench contract)."""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rust_bindings.camera import CameraParameters  # noqa: E402
from tadataka_b200 import _lib, scene  # noqa: E402
from tadataka_b200.mapper import SemiDenseMapper  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=30)
ap.add_argument("--plans", default="2,0")
ap.add_argument("--height", type=int, default=480)
ap.add_argument("--width", type=int, default=640)
args = ap.parse_args()
H, W = args.height, args.width
t0 = time.time()
frames, cam, params = scene.make_sequence(3300, n_frames=args.frames, H=H, W=W)
print(f"rendered {args.frames} frames in {time.time() - t0:.1f}s (numpy)")
cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
depth0 = np.random.default_rng(3300).uniform(params[0], params[1], (H, W))
ref = None
for plan in [int(x) for x in args.plans.split(",")]:
    _lib.check(_lib.set_kernel_plan(plan))
    for rep in range(2):   # the first pass warms allocator and workspaces
        m = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0,
                            fuse=True, regularize=True, prune=True)
        torch.cuda.synchronize()
        t0 = time.time()
        ok = 0
        for img, T_wf, _ in frames[1:]:
            f = m.track(cp, img, T_wf)
        torch.cuda.synchronize()
        dt = time.time() - t0
    maps = m.maps()
    ok = int((maps[3] == 0).sum())
    same = "" if ref is None else f"; maps identical to plan {ref[0]}: {all(np.array_equal(a, b, equal_nan=True) for a, b in zip(ref[1], maps))}"
    if ref is None:
        ref = (plan, maps)
    print(f"plan {plan}: {dt / (args.frames - 1) * 1e3:.3f} ms per frame (wall clock, image upload included), "
          f"{m.n_frames} frames held, last frame: {ok} px updated, max age {int(maps[2].max())}{same}")
_lib.set_kernel_plan(_lib.DEFAULT_PLAN)
