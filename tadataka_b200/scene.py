"""Synthetic piecewise-planar scenes with exact ground truth.

An analytic ray-cast in numpy (no graphics API): one back wall z = Z0 and three slanted textured
quads in front of it, all expressed in world coordinates.  A frame is rendered from any pose, so
a keyframe and its reference frames see the same world and every pixel has an exact depth.
Everything is seeded with numpy.random.default_rng.
"""
import numpy as np

# parameter sets
SET_S = dict(name="S", params=(60.0, 1000.0, 0.01, 0.01, 0.01, 0.2), z0=400.0, quad_z=(120.0, 300.0),
             baseline=10.0, extent=180.0, tex_freq=0.08)                      # Tsukuba-scale
SET_T = dict(name="T", params=(0.1, 10.0, 0.01, 0.01, 0.002, 0.05), z0=4.0, quad_z=(1.2, 3.0),
             baseline=0.05, extent=1.8, tex_freq=8.0)                          # TUM-scale

CAM_VGA = np.array([517.3, 516.5, 318.6, 255.3])     # TUM fr1 (tadataka/dataset/tum_rgbd.py:70-75)
CAM_TSUKUBA = np.array([615.0, 615.0, 320.0, 240.0])  # tadataka/dataset/new_tsukuba.py:98-101
CAM_720P = np.array([900.0, 900.0, 639.5, 359.5])


def rotation(rx, ry, rz):
    """R = Rz Ry Rx, angles in radians."""
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def pose(R=None, t=(0., 0., 0.)):
    T = np.eye(4)
    if R is not None:
        T[:3, :3] = R
    T[:3, 3] = t
    return T


class World:
    def __init__(self, pset, seed):
        rng = np.random.default_rng(seed)
        self.pset = pset
        self.z0 = pset["z0"]
        zlo, zhi = pset["quad_z"]
        ext = pset["extent"]
        self.planes = []   # (p0, normal, a, b, half_a, half_b)
        # back wall: infinite
        self.planes.append((np.array([0., 0., self.z0]), np.array([0., 0., 1.]),
                            np.array([1., 0., 0.]), np.array([0., 1., 0.]), np.inf, np.inf))
        for _ in range(3):
            c = np.array([rng.uniform(-0.5, 0.5) * ext, rng.uniform(-0.4, 0.4) * ext,
                          rng.uniform(zlo, zhi)])
            R = rotation(rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 0.5), rng.uniform(-0.3, 0.3))
            self.planes.append((c, R[:, 2], R[:, 0], R[:, 1], rng.uniform(0.25, 0.5) * ext,
                                rng.uniform(0.25, 0.5) * ext))
        f = pset["tex_freq"]
        self.tex = [(f * rng.uniform(0.7, 1.4), f * rng.uniform(0.7, 1.4), rng.uniform(0, 6.28),
                     rng.uniform(0, 6.28), f * rng.uniform(1.5, 3.0)) for _ in self.planes]
        self.lattice = rng.random((len(self.planes), 64, 64))

    def _noise(self, k, s, t):
        lat = self.lattice[k]
        s0, t0 = np.floor(s), np.floor(t)
        fs, ft = s - s0, t - t0
        fs, ft = fs * fs * (3 - 2 * fs), ft * ft * (3 - 2 * ft)
        i0, j0 = s0.astype(np.int64) % 64, t0.astype(np.int64) % 64
        i1, j1 = (i0 + 1) % 64, (j0 + 1) % 64
        return (lat[j0, i0] * (1 - fs) * (1 - ft) + lat[j0, i1] * fs * (1 - ft)
                + lat[j1, i0] * (1 - fs) * ft + lat[j1, i1] * fs * ft)

    def render(self, T_wc, cam, H, W):
        """-> (image f64 [H,W] in [0,1], depth f64 [H,W] = z in the camera frame)."""
        fx, fy, ox, oy = cam
        xs = (np.arange(W, dtype=np.float64) - ox) / fx
        ys = (np.arange(H, dtype=np.float64) - oy) / fy
        X, Y = np.meshgrid(xs, ys)
        dirs_c = np.stack([X, Y, np.ones_like(X)], axis=-1)           # depth == ray parameter
        R, o = T_wc[:3, :3], T_wc[:3, 3]
        dirs = dirs_c @ R.T
        best_t = np.full((H, W), np.inf)
        image = np.zeros((H, W))
        for k, (p0, n, a, b, ha, hb) in enumerate(self.planes):
            denom = dirs @ n
            with np.errstate(divide="ignore", invalid="ignore"):
                t = ((p0 - o) @ n) / denom
            hit = o + dirs * t[..., None]
            rel = hit - p0
            s, tt = rel @ a, rel @ b
            ok = (t > 1e-9) & np.isfinite(t) & (np.abs(s) <= ha) & (np.abs(tt) <= hb) & (t < best_t)
            fa, fb, p1, p2, fn = self.tex[k]
            # textured islands (contrast mask from low-frequency noise) on a nearly flat ground, so
            # that, as in natural images, most pixels fail the gradient test
            mask = np.clip((self._noise(k, s * fn / 12. + 7.3, tt * fn / 12. + 1.9) - 0.52) * 8., 0., 1.)
            fine = (self._noise(k, s * fn, tt * fn) - 0.5) * 2.0
            coarse = (self._noise(k, s * fn / 4. + 3.1, tt * fn / 4. + 5.7) - 0.5) * 2.0
            wave = np.sin(2 * np.pi * fa * s + p1) * np.sin(2 * np.pi * fb * tt + p2)
            val = 0.5 + mask * (0.35 * fine + 0.1 * wave) + 0.08 * coarse
            image = np.where(ok, np.clip(val, 0., 1.), image)
            best_t = np.where(ok, t, best_t)
        depth = np.where(np.isfinite(best_t), best_t, self.z0)
        return np.ascontiguousarray(image), np.ascontiguousarray(depth)


def make_pair(seed, H=480, W=640, cam=None, pset=SET_S, yaw_deg=1.0):
    """C1-style input: one keyframe, one reference (+x baseline, small yaw), age == 1, prior depth
    = GT * (1 + U(-0.1, 0.1)), prior variance log-uniform in [1e-3, 1]."""
    return make_keyframe(seed, R=1, H=H, W=W, cam=cam, pset=pset, yaw_deg=yaw_deg, ages="ones")


def make_keyframe(seed, R=8, H=480, W=640, cam=None, pset=SET_S, yaw_deg=1.0, ages="blocks"):
    """C2-style input: one keyframe and R references on a seeded random walk.

    ages: "ones" (every pixel uses the newest reference) or "blocks" (seeded integers 0..R in
    32x32 blocks; 0 -> NotProcessed).
    Returns a dict of numpy arrays shaped for ONE keyframe:
      key_img [H,W], key_cam [4], T_wk [4,4], ref_imgs [R,H,W], ref_cams [R,4], T_wr [R,4,4],
      age [H,W] u64, prior_depth, prior_variance, gt_depth [H,W], params (6,)
    """
    cam = CAM_VGA if cam is None else np.asarray(cam, dtype=np.float64)
    rng = np.random.default_rng(seed)
    world = World(pset, seed)
    T_wk = pose()
    key_img, gt = world.render(T_wk, cam, H, W)
    base = pset["baseline"]
    ref_imgs, T_wrs = [], []
    for r in range(R):
        # random walk: mostly sideways, growing baseline for older references
        scale = base * (0.2 + 1.4 * (R - r) / max(R, 1)) if R > 1 else base
        t = np.array([scale * rng.choice([-1., 1.]) * rng.uniform(0.8, 1.2),
                      scale * rng.uniform(-0.15, 0.15), scale * rng.uniform(-0.1, 0.1)])
        Rm = rotation(np.deg2rad(rng.uniform(-0.3, 0.3)), np.deg2rad(yaw_deg * rng.uniform(-1, 1)),
                      np.deg2rad(rng.uniform(-0.2, 0.2)))
        T_wr = pose(Rm, t)
        img, _ = world.render(T_wr, cam, H, W)
        ref_imgs.append(img)
        T_wrs.append(T_wr)
    if ages == "ones":
        age = np.ones((H, W), dtype=np.uint64)
    else:
        by, bx = (H + 31) // 32, (W + 31) // 32
        blocks = rng.integers(0, R + 1, size=(by, bx))
        age = np.kron(blocks, np.ones((32, 32), dtype=np.int64))[:H, :W].astype(np.uint64)
    prior_depth = gt * (1.0 + rng.uniform(-0.1, 0.1, size=(H, W)))
    lo, hi = pset["params"][0], pset["params"][1]
    # the inverse-depth "variance" is used linearly as a half-width; 1e-3 .. 1 spans
    # narrow searches up to the full valid range (tests/vo/semi_dense/test_semi_dense.py:64-65)
    prior_variance = 10.0 ** rng.uniform(-3.0, 0.0, size=(H, W))
    prior_depth = np.clip(prior_depth, lo * 1.01, hi * 0.99)
    return dict(key_img=key_img, key_cam=cam.copy(), T_wk=T_wk, ref_imgs=np.stack(ref_imgs),
                ref_cams=np.tile(cam, (R, 1)), T_wr=np.stack(T_wrs), age=np.ascontiguousarray(age),
                prior_depth=np.ascontiguousarray(prior_depth),
                prior_variance=np.ascontiguousarray(prior_variance), gt_depth=gt,
                params=np.array(pset["params"], dtype=np.float64))


def stack_keyframes(items):
    """list of make_keyframe() dicts (same R, H, W) -> batched arrays [P, ...]."""
    keys = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth",
            "prior_variance", "gt_depth"]
    out = {k: np.ascontiguousarray(np.stack([it[k] for it in items])) for k in keys}
    out["params"] = items[0]["params"]
    return out


def sequence_poses(seed, n_frames=30, pset=SET_S):
    """The smooth seeded trajectory of make_sequence: n_frames camera-to-world poses [n,4,4]."""
    rng = np.random.default_rng(seed)
    base = pset["baseline"]
    vel = np.array([base * 0.5, base * 0.05, base * 0.15])
    phase = rng.uniform(0, 6.28, size=3)
    poses = []
    for i in range(n_frames):
        t = vel * i + 0.2 * base * np.sin(0.3 * i + phase)
        Rm = rotation(np.deg2rad(0.3 * np.sin(0.2 * i + phase[0])), np.deg2rad(0.15 * i * 0.3),
                      np.deg2rad(0.2 * np.sin(0.25 * i + phase[2])))
        poses.append(pose(Rm, t))
    return np.stack(poses)


def make_sequence(seed, n_frames=30, H=480, W=640, cam=None, pset=SET_S):
    """C3-style input: frames on a smooth seeded trajectory with ground-truth poses.
    -> list of (image, T_wf, gt_depth), camera vector, params."""
    cam = CAM_VGA if cam is None else np.asarray(cam, dtype=np.float64)
    world = World(pset, seed)
    frames = []
    for T_wf in sequence_poses(seed, n_frames, pset):
        img, gt = world.render(T_wf, cam, H, W)
        frames.append((img, T_wf, gt))
    return frames, cam.copy(), np.array(pset["params"], dtype=np.float64)
