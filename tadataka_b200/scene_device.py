"""Synthetic scenes rendered on the GPU (SURVEY.md 8d workload, 8f rank 3).

`render` is the device version of scene.World.render (csrc/scene.cu, tdk_render_scene_dev);
`make_keyframes` builds a whole batch of C2-style keyframes -- images by the ray-caster, age
blocks and priors by torch's device generator -- without rendering anything in numpy or moving an
image through PCIe.  The reference has no counterpart: it reads frames from datasets.
"""
import numpy as np
import torch

from . import _lib, scene


def world_tensors(world, device):
    """scene.World -> (planes [K,14], tex [K,5], lattice [K,64,64]) as f64 tensors on `device`."""
    planes = np.array([np.concatenate([p0, n, a, b, [ha, hb]]) for p0, n, a, b, ha, hb in world.planes])
    tex = np.array(world.tex, dtype=np.float64)
    to = lambda x: torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).to(device)
    return to(planes), to(tex), to(world.lattice)


def render(world, T_wc, cams, H, W, device="cuda"):
    """N views of one world.  T_wc [N,4,4], cams [N,4] (numpy or tensors) ->
    (image [N,H,W], depth [N,H,W]) f64 tensors on `device`."""
    dev = torch.device(device)
    planes, tex, lattice = world_tensors(world, dev)
    T = torch.as_tensor(np.asarray(T_wc), dtype=torch.float64).reshape(-1, 16).contiguous().to(dev)
    c = torch.as_tensor(np.asarray(cams), dtype=torch.float64).reshape(-1, 4).contiguous().to(dev)
    N = T.shape[0]
    if c.shape[0] != N:
        raise ValueError(f"{N} poses but {c.shape[0]} cameras")
    image = torch.empty((N, H, W), dtype=torch.float64, device=dev)
    depth = torch.empty((N, H, W), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.render_scene_dev(planes.data_ptr(), tex.data_ptr(), lattice.data_ptr(), planes.shape[0],
                                         float(world.z0), T.data_ptr(), c.data_ptr(), N, H, W,
                                         image.data_ptr(), depth.data_ptr(),
                                         torch.cuda.current_stream(dev).cuda_stream))
    return image, depth


def make_keyframes(seeds, R=8, H=480, W=640, cam=None, pset=scene.SET_S, yaw_deg=1.0, device="cuda",
                   ages="blocks", per_item_rng=False):
    """One C2-style keyframe per seed, every one its own world: the batched arrays of
    scene.stack_keyframes as device tensors (same keys, plus params as a numpy array).  Poses follow
    scene.make_keyframe's seeded random walk; age blocks and priors come from a device generator
    seeded with seeds[0], so the batch is reproducible but not equal to the numpy one.
    ages="ones": every pixel uses the newest reference (C1 / C5 pairs).  per_item_rng=True: the
    generator is re-seeded with each keyframe's own seed, so a keyframe does not depend on the batch
    it is built in (what a sharded run needs to be comparable with a single-process one)."""
    dev = torch.device(device)
    cam = scene.CAM_VGA if cam is None else np.asarray(cam, dtype=np.float64)
    P = len(seeds)
    out = {k: [] for k in ("key_img", "ref_imgs", "gt_depth", "T_wk", "T_wr")}
    base = pset["baseline"]
    for seed in seeds:
        rng = np.random.default_rng(seed)
        world = scene.World(pset, seed)
        poses = [scene.pose()]
        for r in range(R):
            scale = base * (0.2 + 1.4 * (R - r) / max(R, 1)) if R > 1 else base
            t = np.array([scale * rng.choice([-1., 1.]) * rng.uniform(0.8, 1.2),
                          scale * rng.uniform(-0.15, 0.15), scale * rng.uniform(-0.1, 0.1)])
            Rm = scene.rotation(np.deg2rad(rng.uniform(-0.3, 0.3)), np.deg2rad(yaw_deg * rng.uniform(-1, 1)),
                                np.deg2rad(rng.uniform(-0.2, 0.2)))
            poses.append(scene.pose(Rm, t))
        poses = np.stack(poses)
        img, dep = render(world, poses, np.tile(cam, (R + 1, 1)), H, W, dev)
        out["key_img"].append(img[0]); out["ref_imgs"].append(img[1:]); out["gt_depth"].append(dep[0])
        out["T_wk"].append(torch.from_numpy(poses[0]).to(dev)); out["T_wr"].append(torch.from_numpy(poses[1:]).to(dev))
    b = {k: torch.stack(v).contiguous() for k, v in out.items()}
    g = torch.Generator(device=dev)
    by, bx = (H + 31) // 32, (W + 31) // 32

    def draw(n):
        blocks = torch.randint(0, R + 1, (n, by, bx), generator=g, device=dev)
        noise = torch.rand((n, H, W), generator=g, device=dev, dtype=torch.float64) * 0.2 - 0.1
        logv = torch.rand((n, H, W), generator=g, device=dev, dtype=torch.float64) * 3.0 - 3.0
        return blocks, noise, logv

    if per_item_rng:
        parts = []
        for seed in seeds:
            g.manual_seed(int(seed))
            parts.append(draw(1))
        blocks, noise, logv = (torch.cat([q[i] for q in parts]) for i in range(3))
    else:
        g.manual_seed(int(seeds[0]))
        blocks, noise, logv = draw(P)
    if ages == "ones":
        b["age"] = torch.ones((P, H, W), dtype=torch.int64, device=dev)
    else:
        b["age"] = blocks.repeat_interleave(32, 1).repeat_interleave(32, 2)[:, :H, :W].contiguous().to(torch.int64)
    lo, hi = pset["params"][0], pset["params"][1]
    b["prior_depth"] = torch.clamp(b["gt_depth"] * (1.0 + noise), lo * 1.01, hi * 0.99).contiguous()
    b["prior_variance"] = (10.0 ** logv).contiguous()
    camt = torch.from_numpy(cam).to(dev)
    b["key_cam"] = camt.repeat(P, 1).contiguous()
    b["ref_cams"] = camt.repeat(P, R, 1).contiguous()
    b["params"] = np.array(pset["params"], dtype=np.float64)
    return b
