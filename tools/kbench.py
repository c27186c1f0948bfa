"""Quick kernel-only timing + exact parity check of update_depth on the C2 workload shape.
    python tools/kbench.py [--keyframes 32] [--distinct 4] [--iters 20] [--lanes 8,16,32] [--c4]
Development tool (not part of the bench contract)."""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import oracle as ora  # noqa: E402
from tadataka_b200 import _lib, device, scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--keyframes", type=int, default=32)
ap.add_argument("--distinct", type=int, default=4)
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--plans", default="0,1")
ap.add_argument("--c4", action="store_true", help="720p TUM-scale pairs instead of C2")
ap.add_argument("--no-parity", action="store_true")
ap.add_argument("--device-scenes", action="store_true", help="every keyframe its own world from the CUDA ray-caster (fast set-up; implies --no-parity)")
ap.add_argument("--smooth", action="store_true", help="prior variance constant per 32x32 block, prior depth = ground truth with a smooth 3 %% bias (what a propagated map looks like) instead of per-pixel noise")
args = ap.parse_args()

if args.device_scenes:
    from tadataka_b200 import scene_device
    args.no_parity = True
    if args.c4:
        tt = scene_device.make_keyframes([4000 + i for i in range(args.keyframes)], R=1, H=720, W=1280, cam=scene.CAM_720P,
                                         pset=scene.SET_T, ages="ones", per_item_rng=True)
    else:
        tt = scene_device.make_keyframes([2000 + i for i in range(args.keyframes)], R=8)
    b = {"params": tt["params"], "key_img": tt["key_img"]}
elif args.c4:
    b = scene.stack_keyframes([scene.make_pair(4000 + i, H=720, W=1280, cam=scene.CAM_720P, pset=scene.SET_T)
                               for i in range(args.keyframes)])
else:
    b = bench.make_workload(0, args.keyframes, args.distinct)
P, H, W = b["key_img"].shape
if args.smooth:
    rng = np.random.default_rng(11)
    by, bx = (H + 31) // 32, (W + 31) // 32
    blocks = 10.0 ** rng.uniform(-3.0, 0.0, size=(P, by, bx))
    b["prior_variance"] = np.ascontiguousarray(np.kron(blocks, np.ones((32, 32)))[:, :H, :W])
    yy, xx = np.mgrid[0:H, 0:W]
    bias = 1.0 + 0.03 * np.sin(xx / 97.0) * np.cos(yy / 71.0)
    lo, hi = b["params"][0], b["params"][1]
    b["prior_depth"] = np.ascontiguousarray(np.clip(b["gt_depth"] * bias, lo * 1.01, hi * 0.99))
names = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth", "prior_variance"]
t = {k: tt[k] for k in names} if args.device_scenes else {k: torch.from_numpy(b[k]).cuda() for k in names}
if not args.no_parity:
    t0 = time.time()
    od, ov, of, oam, onn = ora.update_depth_batch(*(b[k] for k in names), ora.make_params(*b["params"]),
                                                  threads=bench.host_threads(), debug=True)
    print(f"oracle {time.time() - t0:.2f}s; searched {(oam >= 0).mean() * 100:.1f}% success {(of == 0).mean() * 100:.1f}% "
          f"n median {np.median(onn[oam >= 0]):.0f} max {onn.max()}")
for lanes in [int(x) for x in args.plans.split(",")]:
    _lib.check(_lib.set_kernel_plan(lanes))
    d, v, f, am = device.update_depth(*(t[k] for k in names), b["params"], debug=True)
    torch.cuda.synchronize()
    if not args.no_parity:
        ok = (np.array_equal(d.cpu().numpy(), od, equal_nan=True) and np.array_equal(v.cpu().numpy(), ov, equal_nan=True)
              and np.array_equal(f.cpu().numpy(), of) and np.array_equal(am.cpu().numpy(), oam))
        print(f"plan {lanes}: parity {'EXACT' if ok else 'MISMATCH'}", end="; ")
        if not ok:
            print("flag agree", (f.cpu().numpy() == of).mean(), "argmin agree", (am.cpu().numpy() == oam).mean())
    out = (d, v, f)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        device.update_depth(*(t[k] for k in names), b["params"], check=False, out=out, fuse=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.iters):
        device.update_depth(*(t[k] for k in names), b["params"], check=False, out=out, fuse=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.iters
    print(f"{ms * 1e3 / P:.2f} us/keyframe  {P * H * W / ms / 1e6:.2f} Gpx/s  roofline {64 * P * H * W / ms / 1e6 / 6555.8 * 100:.1f}%")
