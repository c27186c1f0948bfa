"""Device-resident keyframe state for the semi-dense mapping loop.

The reference's caller (examples/semi_dense_vo.py:152-207) keeps the depth / variance / age maps
and the list of reference frames in numpy and hands them to the Rust extension frame by frame:
    age1 = increment_age(age0, cam0, cam1, T10, depth0)
    depth1, var1 = propagate(T10, cam0, cam1, depth0, var0, default_depth, default_variance, bias)
    depth1, var1, flag = update_depth(frame1, refframes, age1, depth1, var1, params)
    refframes.append(frame1)
`SemiDenseMapper` runs the same sequence through the native mapper of libtdk_b200.so
(csrc/mapper.cu, tdk_mapper_*): everything stays in one device arena, a new frame costs ONE C-ABI
call and one image upload (H*W*8 bytes as f64; H*W as 8-bit gray, converted on the device), and
track() does not wait for the GPU.  Optionally each successful measurement is fused with the
propagated prior (Gaussian fusion in inverse depth, fusion.rs:3-11) and the result smoothed by the
3x3 regulariser (regularization.rs:29-64), as the (dead) update_hypothesis of
semi_dense_vo.py:82-88 intended.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib, device
from .camera import camera_vector


def _image_format(image):
    """numpy frame -> (contiguous array, TDK_IMAGE_* code)."""
    image = np.ascontiguousarray(image)
    if image.dtype == np.uint8:
        if image.ndim == 2:
            return image, _lib.IMAGE_U8
        if image.ndim == 3 and image.shape[2] == 3:
            return image, _lib.IMAGE_RGB8
        if image.ndim == 3 and image.shape[2] == 4:
            return image, _lib.IMAGE_RGBA8
        raise ValueError("8-bit frames must be [H,W], [H,W,3] or [H,W,4]")
    if image.ndim != 2:
        raise ValueError("a float frame must be a 2-D gray image")
    return np.ascontiguousarray(image, dtype=np.float64), _lib.IMAGE_F64


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class SemiDenseMapper:
    def __init__(self, camera_params, image, transform_wf, params, default_depth, default_variance,
                 uncertaintity_bias, depth_map=None, variance_map=None, fuse=False, regularize=False,
                 prune=False, capacity=64):
        """First frame: age 0 everywhere; depth defaults to `default_depth`, variance to
        `default_variance` (semi_dense_vo.py:210-219 draws the depth at random instead: pass it).
        `image`: float64 [H,W] gray, or uint8 [H,W] / [H,W,3] / [H,W,4] (converted on the device).
        `capacity`: frames the ring of reference images holds; with prune=True frames no pixel's age
        points to any more are dropped, so the ring only has to cover the oldest age on the map."""
        device._require_cuda()
        image, fmt = _image_format(image)
        H, W = image.shape[:2]
        self.shape = (H, W)
        self.params = device.params_array(params)
        self.fuse, self.smooth, self.prune = bool(fuse), bool(regularize), bool(prune)
        self._dev = torch.device("cuda", torch.cuda.current_device())
        nbytes = _lib.mapper_arena_bytes(H, W, int(capacity))
        self._arena = torch.empty(nbytes, dtype=torch.uint8, device=self._dev)
        handle = C.c_void_p()
        _lib.check(_lib.mapper_create(C.c_void_p(self._arena.data_ptr()), nbytes, H, W, int(capacity), _p(self.params),
                                      float(default_depth), float(default_variance), float(uncertaintity_bias),
                                      int(self.fuse), int(self.smooth), int(self.prune),
                                      C.c_void_p(torch.cuda.current_stream(self._dev).cuda_stream), C.byref(handle)))
        self._h = handle
        self._flag_view = None
        self._last_pose = np.ascontiguousarray(transform_wf, dtype=np.float64)
        dm = None if depth_map is None else np.ascontiguousarray(depth_map, dtype=np.float64)
        vm = None if variance_map is None else np.ascontiguousarray(variance_map, dtype=np.float64)
        for a in (dm, vm):
            if a is not None and a.shape != (H, W):
                raise ValueError("depth_map / variance_map must have the frame's shape")
        _lib.check(_lib.mapper_reset(self._h, _p(image), fmt, _p(camera_vector(camera_params)), _p(self._last_pose),
                                     _p(dm), _p(vm)))

    def __del__(self):
        h = getattr(self, "_h", None)
        destroy = getattr(_lib, "mapper_destroy", None)   # may be gone already at interpreter shutdown
        if h and destroy is not None:
            destroy(h)
            self._h = None

    # ---- state
    def _state(self):
        n, ring_start = C.c_int64(), C.c_int64()
        offs = [C.c_size_t() for _ in range(5)]
        _lib.check(_lib.mapper_state(self._h, C.byref(n), *[C.byref(o) for o in offs], C.byref(ring_start)))
        return n.value, [o.value for o in offs], ring_start.value

    def _view(self, off, dtype):
        H, W = self.shape
        return self._arena[off:off + H * W * 8].view(dtype).view(H, W)

    @property
    def n_frames(self):
        return self._state()[0]

    @property
    def depth(self):
        return self._view(self._state()[1][0], torch.float64)

    @property
    def variance(self):
        return self._view(self._state()[1][1], torch.float64)

    @property
    def age(self):
        return self._view(self._state()[1][2], torch.uint64)

    @age.setter
    def age(self, value):
        """Overwrite the current age map (e.g. a re-initialisation) in place."""
        self._view(self._state()[1][2], torch.uint64).copy_(value)

    @property
    def flag(self):
        if self._flag_view is None:     # the flag map never changes its place in the arena
            self._flag_view = self._view(self._state()[1][3], torch.int64)
        return self._flag_view

    def reference_images(self):
        """The frames currently held (device tensor [n_frames, H, W], a view of the ring)."""
        n, offs, start = self._state()
        H, W = self.shape
        hw8 = H * W * 8
        a = offs[4] + start * hw8
        return self._arena[a:a + n * hw8].view(torch.float64).view(n, H, W)

    # ---- the loop
    def track(self, camera_params, image, transform_wf):
        """Advance to a new frame with a known pose (world <- frame).  Returns the flag map (a device
        tensor that the next track() overwrites); the maps are in .depth / .variance / .age.  Does
        not wait for the GPU: an error of this frame's update (an age beyond the ring, a singular
        pose) is raised by one of the next two track() calls, by maps() or by sync()."""
        image, fmt = _image_format(image)
        if image.shape[:2] != self.shape:
            raise ValueError("every frame must have the first frame's shape")
        T_w1 = np.ascontiguousarray(transform_wf, dtype=np.float64)
        T10 = np.ascontiguousarray(np.linalg.inv(T_w1) @ self._last_pose)   # semi_dense_vo.py: transform10
        _lib.check(_lib.mapper_track(self._h, _p(image), fmt, _p(camera_vector(camera_params)), _p(T_w1), _p(T10)))
        self._last_pose = T_w1
        return self.flag

    def sync(self):
        _lib.check(_lib.mapper_sync(self._h))

    def _drop_unused_frames(self):
        """Census of the current age map and pruning in one step (waits for the device)."""
        _lib.check(_lib.mapper_prune(self._h))

    def maps(self):
        """(depth, variance, age, flag) as numpy arrays (one device -> host copy each)."""
        self.sync()
        return (self.depth.cpu().numpy(), self.variance.cpu().numpy(), self.age.cpu().numpy(),
                self.flag.cpu().numpy())
