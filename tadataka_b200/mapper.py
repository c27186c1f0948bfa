"""Device-resident keyframe state for the semi-dense mapping loop.

The reference's caller (examples/semi_dense_vo.py:152-207) keeps the depth / variance / age maps
and the list of reference frames in numpy and hands them to the Rust extension frame by frame:
    age1 = increment_age(age0, cam0, cam1, T10, depth0)
    depth1, var1 = propagate(T10, cam0, cam1, depth0, var0, default_depth, default_variance, bias)
    depth1, var1, flag = update_depth(frame1, refframes, age1, depth1, var1, params)
    refframes.append(frame1)
`SemiDenseMapper` runs the same sequence with everything resident in HBM: a new frame costs one
image upload (H*W*8 bytes); the maps and the reference images never cross PCIe until asked for.
Optionally each successful measurement is fused with the propagated prior (Gaussian fusion in
inverse depth, fusion.rs:3-11) and the result smoothed by the 3x3 regulariser
(regularization.rs:29-64), as the (dead) update_hypothesis of semi_dense_vo.py:82-88 intended.
"""
import numpy as np
import torch

from . import device
from .camera import camera_vector


def _cu(a, dtype=np.float64):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=dtype)).cuda()


def _upload(a):
    """numpy f64 array -> device tensor through pinned memory (torch's caching host allocator), so
    the copy is asynchronous and the host goes on preparing the next launch."""
    src = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
    pinned = torch.empty(src.shape, dtype=src.dtype, pin_memory=True)
    pinned.copy_(src)
    return pinned.cuda(non_blocking=True)


class SemiDenseMapper:
    def __init__(self, camera_params, image, transform_wf, params, default_depth, default_variance,
                 uncertaintity_bias, depth_map=None, variance_map=None, fuse=False, regularize=False,
                 prune=False):
        """First frame: age 0 everywhere; depth defaults to `default_depth`, variance to
        `default_variance` (semi_dense_vo.py:210-219 draws the depth at random instead: pass it)."""
        device._require_cuda()
        image = np.ascontiguousarray(image, dtype=np.float64)
        H, W = image.shape
        self.shape = (H, W)
        self.params = device.params_array(params)
        self.default_depth, self.default_variance = float(default_depth), float(default_variance)
        self.bias = float(uncertaintity_bias)
        self.fuse, self.smooth = bool(fuse), bool(regularize)
        self.prune = bool(prune)   # drop reference frames no pixel's age points to any more
        self.depth = _cu(np.full((H, W), self.default_depth) if depth_map is None else depth_map)
        self.variance = _cu(np.full((H, W), self.default_variance) if variance_map is None
                            else variance_map)
        self.age = torch.zeros((H, W), dtype=torch.uint64, device="cuda")
        self.flag = torch.full((H, W), -9, dtype=torch.int64, device="cuda")
        self._images = [_cu(image)]
        self._cams = [camera_vector(camera_params)]
        self._poses = [np.ascontiguousarray(transform_wf, dtype=np.float64)]
        self._pending = None   # age census of the last map, on its way to the host

    @property
    def n_frames(self):
        return len(self._images)

    def track(self, camera_params, image, transform_wf):
        """Advance to a new frame with a known pose (world <- frame).  Returns the flag map
        (device tensor); the maps are in .depth / .variance / .age."""
        H, W = self.shape
        image = np.ascontiguousarray(image, dtype=np.float64)
        if image.shape != (H, W):
            raise ValueError("every frame must have the first frame's shape")
        cam1 = camera_vector(camera_params)
        T_w1 = np.ascontiguousarray(transform_wf, dtype=np.float64)
        self._apply_pending_prune()
        cam0, T_w0 = self._cams[-1], self._poses[-1]
        T10 = np.linalg.inv(T_w1) @ T_w0            # semi_dense_vo.py: transform10
        R = len(self._images)
        # one pinned staging buffer and one copy for the frame's small operands, one for the image
        small = np.concatenate([T10.ravel(), cam0, cam1, T_w1.ravel(), np.concatenate(self._cams),
                                np.concatenate([p.ravel() for p in self._poses])])
        g_small = _upload(small)
        g_T10, g_cam0, g_cam1 = g_small[0:16].view(4, 4), g_small[16:20], g_small[20:24]
        g_Tw1 = g_small[24:40].view(4, 4)
        g_cams = g_small[40:40 + 4 * R].view(R, 4)
        g_poses = g_small[40 + 4 * R:40 + 20 * R].view(R, 4, 4)
        g_img = _upload(image)
        # increment_age + propagate (semi_dense_vo.py:185-194) through one warp of the map
        depth1, var1, age1 = device.propagate_with_age(g_T10, g_cam0, g_cam1, self.depth, self.variance,
                                                       self.age, self.default_depth, self.default_variance,
                                                       self.bias)
        d, v, f = device.update_depth_frames(g_img, g_cam1, g_Tw1, self._images, g_cams, g_poses,
                                             age1, depth1, var1, self.params, fuse=self.fuse)
        if self.smooth:
            d = device.regularize(d, v, f)
        self.depth, self.variance, self.age, self.flag = d, v, age1, f
        self._images.append(g_img)
        self._cams.append(cam1)
        self._poses.append(T_w1)
        assert len(self._images) == R + 1
        if self.prune:
            self._count_ages()
        return f

    def _count_ages(self):
        """Launch the age census of the new map and start its copy to pinned host memory; the
        pruning it allows is applied at the start of the next track(), so no frame waits for it."""
        n = len(self._images)
        counts = device.age_usage(self.age, n)
        host = torch.empty(counts.shape, dtype=counts.dtype, pin_memory=True)
        host.copy_(counts, non_blocking=True)
        done = torch.cuda.Event()
        done.record()
        self._pending = (host, done, n)

    def _apply_pending_prune(self):
        if self._pending is None:
            return
        host, done, n = self._pending
        self._pending = None
        done.synchronize()
        self._drop_frames_older_than(host.numpy(), n)

    def _drop_unused_frames(self):
        """Census and pruning in one step (waits for the device)."""
        n = len(self._images)
        self._pending = None
        self._drop_frames_older_than(device.age_usage(self.age, n).cpu().numpy(), n)

    def _drop_frames_older_than(self, counts, n):
        """The reference's TODO (examples/semi_dense_vo.py:202).  update_depth reads
        refframes[len - age], so the frames older than the largest age on the map are dead; dropping
        them from the front leaves len - age pointing at the same frames.  `counts` = pixels per age
        for the map that went with a ring of n frames (frames appended since then are newer and stay).
        The newest frame always stays (the next increment_age needs its pose and camera)."""
        used = np.nonzero(counts)[0]
        oldest = int(used.max()) if used.size else 0
        # at the next update the ages have grown by one and the ring by one: frame index
        # (n + 1) - 1 - (oldest + 1) = n - 1 - oldest is the oldest one still read
        keep_from = max(0, min(n - 1 - oldest, n - 1))
        if keep_from > 0:
            del self._images[:keep_from], self._cams[:keep_from], self._poses[:keep_from]

    def maps(self):
        """(depth, variance, age, flag) as numpy arrays (one device -> host copy each)."""
        return (self.depth.cpu().numpy(), self.variance.cpu().numpy(), self.age.cpu().numpy(),
                self.flag.cpu().numpy())
