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
        cam0, T_w0 = self._cams[-1], self._poses[-1]
        T10 = np.linalg.inv(T_w1) @ T_w0            # semi_dense_vo.py: transform10
        g_T10, g_cam0, g_cam1 = _cu(T10), _cu(cam0), _cu(cam1)
        age1 = device.increment_age(self.age, g_cam0, g_cam1, g_T10, self.depth)
        depth1, var1 = device.propagate(g_T10, g_cam0, g_cam1, self.depth, self.variance,
                                        self.default_depth, self.default_variance, self.bias)
        g_img = _cu(image)
        R = len(self._images)
        d, v, f = device.update_depth_frames(g_img, g_cam1, _cu(T_w1), self._images,
                                             _cu(np.stack(self._cams)), _cu(np.stack(self._poses)),
                                             age1, depth1, var1, self.params, fuse=self.fuse)
        if self.smooth:
            d = device.regularize(d, v, f)
        self.depth, self.variance, self.age, self.flag = d, v, age1, f
        self._images.append(g_img)
        self._cams.append(cam1)
        self._poses.append(T_w1)
        assert len(self._images) == R + 1
        if self.prune:
            self._drop_unused_frames()
        return f

    def _drop_unused_frames(self):
        """The reference's TODO (examples/semi_dense_vo.py:202).  update_depth reads
        refframes[len - age], so the frames older than the largest age on the map are dead; dropping
        them from the front leaves len - age pointing at the same frames.  The current frame always
        stays (the next increment_age needs its pose and camera)."""
        n = len(self._images)
        counts = device.age_usage(self.age, n).cpu().numpy()   # ages 0..n, then older (impossible)
        used = np.nonzero(counts)[0]
        oldest = int(used.max()) if used.size else 0
        # at the next track() the ages have grown by one and the list by one: frame index
        # (n + 1) - (oldest + 1) = n - oldest is the oldest one still read
        keep_from = max(0, min(n - 1 - oldest, n - 1))
        if keep_from > 0:
            del self._images[:keep_from], self._cams[:keep_from], self._poses[:keep_from]

    def maps(self):
        """(depth, variance, age, flag) as numpy arrays (one device -> host copy each)."""
        return (self.depth.cpu().numpy(), self.variance.cpu().numpy(), self.age.cpu().numpy(),
                self.flag.cpu().numpy())
