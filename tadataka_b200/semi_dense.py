"""Drop-in for the reference's `rust_bindings.semi_dense` extension module
(src/py/semi_dense.rs:235-246): `Frame`, `Params`, `update_depth`, `propagate`, `increment_age`,
`estimate_debug_`, same argument order, dtypes and return shapes, numpy in / numpy out.
`fusion` and `regularize` are exported in addition (the reference keeps them internal or
disabled: fusion.rs, regularization.rs; call shapes from examples/semi_dense_depth.py:41-42 and
src/py/semi_dense.rs:220-233).

Everything is computed by libtdk_b200.so on the current CUDA device; there is no CPU path.

Deviations from the reference, all at the error boundary:
  * shape mismatches raise ValueError (the reference panics: semi_dense.rs:168-173);
  * an age larger than len(refframes) raises ValueError (the reference prints and calls
    std::process::exit(1): semi_dense.rs:202-205);
  * a singular reference pose raises ValueError (the reference panics in inv().unwrap()).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib, device
from .camera import CameraParameters, camera_vector
from .flag import ResultFlag  # noqa: F401  (re-exported for convenience)


def _need(a, name, dtype, ndim):
    if not isinstance(a, np.ndarray) or a.dtype != dtype or a.ndim != ndim:
        raise TypeError(f"{name} must be a {ndim}-D numpy array of dtype {np.dtype(dtype).name}")
    return np.ascontiguousarray(a)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _cuda(a):
    """numpy -> CUDA tensor on the current device (pageable H2D copy)."""
    device._require_cuda()
    return torch.from_numpy(a).cuda()


class Frame:
    """`Frame(camera_params, image, transform)`: image float64 [H,W]; transform float64 [4,4],
    world <- frame.  Image and transform are copied (src/py/semi_dense.rs:53-67); the image is
    additionally uploaded once and kept on the device, so a frame that serves as reference for
    many updates crosses PCIe once."""

    def __init__(self, camera_params, image, transform):
        self._cam = camera_vector(camera_params)
        self._image = _need(image, "image", np.float64, 2).copy()
        self._transform = _need(transform, "transform", np.float64, 2).copy()
        if self._transform.shape != (4, 4):
            raise ValueError("transform must be 4x4")
        self._dev = None

    @property
    def camera_params(self):
        return CameraParameters((self._cam[0], self._cam[1]), (self._cam[2], self._cam[3]))

    @property
    def image(self):
        return self._image.copy()

    @property
    def transform_wf(self):
        return self._transform.copy()

    @property
    def shape(self):
        return self._image.shape

    def device_image(self):
        dev = torch.cuda.current_device()
        if self._dev is None or self._dev.device.index != dev:
            self._dev = _cuda(self._image)
            torch.cuda.current_stream().synchronize()   # the C-ABI call reads it from the library's own stream
        return self._dev


class Params:
    """`Params(min_depth, max_depth, geo_coeff, photo_coeff, ref_step_size, min_gradient)`
    (src/py/semi_dense.rs:91-109): six floats, positional or keyword."""

    def __init__(self, min_depth, max_depth, geo_coeff, photo_coeff, ref_step_size, min_gradient):
        self._v = np.array([min_depth, max_depth, geo_coeff, photo_coeff, ref_step_size,
                            min_gradient], dtype=np.float64)

    def as_array(self):
        return self._v


def _check_frame(f, name):
    if not isinstance(f, Frame):
        raise TypeError(f"{name} must be a Frame")
    return f


def update_depth(keyframe, refframes, age_map, prior_depth, prior_variance, params):
    """-> (depth_map f64, variance_map f64, flag_map int64), fresh arrays.

    Pixel (y, x) is matched against `refframes[len(refframes) - age_map[y, x]]`; age 0 means
    "not processed" (src/semi_dense/semi_dense.rs:190-207).  The returned maps hold the NEW
    measurement where the flag is SUCCESS and the prior elsewhere (no fusion happens here).
    """
    _check_frame(keyframe, "keyframe")
    refframes = list(refframes)
    for i, f in enumerate(refframes):
        _check_frame(f, f"refframes[{i}]")
    age_map = _need(age_map, "age_map", np.uint64, 2)
    prior_depth = _need(prior_depth, "prior_depth", np.float64, 2)
    prior_variance = _need(prior_variance, "prior_variance", np.float64, 2)
    if not isinstance(params, Params):
        raise TypeError("params must be a Params")
    shape = keyframe.shape
    if age_map.shape != shape or prior_depth.shape != shape or prior_variance.shape != shape:
        raise ValueError("age_map, prior_depth and prior_variance must have the keyframe's shape")
    for f in refframes:
        if f.shape != shape:
            raise ValueError("every reference frame must have the keyframe's shape")
    if len(refframes) == 0:
        if age_map.any():
            raise ValueError("Age exceeds the refframe size")
        return (prior_depth.copy(), prior_variance.copy(),
                np.full(shape, int(ResultFlag.NOT_PROCESSED), dtype=np.int64))
    # frames stay on the device (uploaded once per Frame); cameras, poses and the maps go through
    # ONE host-pointer C-ABI call, which moves the numpy maps through the library's pinned ring
    R = len(refframes)
    H, W = shape
    ref_cams = np.ascontiguousarray(np.stack([f._cam for f in refframes]))
    T_wr = np.ascontiguousarray(np.stack([f._transform for f in refframes]))
    key_ptr = (C.c_void_p * 1)(keyframe.device_image().data_ptr())
    ref_ptr = (C.c_void_p * R)(*[f.device_image().data_ptr() for f in refframes])
    d, v = np.empty(shape), np.empty(shape)
    fl = np.empty(shape, dtype=np.int64)
    _lib.check(_lib.update_depth_maps_host(key_ptr, _p(keyframe._cam), _p(keyframe._transform), ref_ptr, _p(ref_cams),
                                           _p(T_wr), _p(age_map), _p(prior_depth), _p(prior_variance),
                                           _p(params.as_array()), 1, R, H, W, _p(d), _p(v), _p(fl), None, None, 0))
    return d, v, fl


def estimate_debug_(u_key, prior_depth, prior_variance, keyframe, refframe, params):
    """Single-pixel probe -> (depth, variance, flag:int); echoes the prior on any failure
    (src/py/semi_dense.rs:126-155).  u_key is an int64 array (x, y)."""
    u_key = _need(u_key, "u_key", np.int64, 1)
    _check_frame(keyframe, "keyframe")
    _check_frame(refframe, "refframe")
    if not isinstance(params, Params):
        raise TypeError("params must be a Params")
    if keyframe.shape != refframe.shape:
        raise ValueError("keyframe and refframe must have the same shape")
    H, W = keyframe.shape
    out = np.zeros(2, dtype=np.float64)
    flag = np.zeros(1, dtype=np.int64)
    device._require_cuda()

    def p(a):
        return a.ctypes.data_as(C.c_void_p)

    rc = _lib.estimate_debug_host(p(u_key), float(prior_depth), float(prior_variance),
                                  p(keyframe._image), p(keyframe._cam), p(keyframe._transform),
                                  p(refframe._image), p(refframe._cam), p(refframe._transform),
                                  p(params.as_array()), H, W, p(out), p(flag), None)
    _lib.check(rc)
    return float(out[0]), float(out[1]), int(flag[0])


def propagate(transform10, camera_params0, camera_params1, depth_map0, variance_map0,
              default_depth, default_variance, uncertaintity_bias):
    """-> (depth_map1, variance_map1).  `transform10` = T_1<-0 (src/py/semi_dense.rs:189-218;
    the parameter really is spelt `uncertaintity_bias`).  One host-pointer C-ABI call
    (tdk_propagate_host)."""
    transform10 = _need(transform10, "transform10", np.float64, 2)
    depth_map0 = _need(depth_map0, "depth_map0", np.float64, 2)
    variance_map0 = _need(variance_map0, "variance_map0", np.float64, 2)
    if transform10.shape != (4, 4) or depth_map0.shape != variance_map0.shape:
        raise ValueError("bad shapes")
    device._require_cuda()
    H, W = depth_map0.shape
    cam0, cam1 = camera_vector(camera_params0), camera_vector(camera_params1)
    d1, v1 = np.empty((H, W)), np.empty((H, W))
    _lib.check(_lib.propagate_host(_p(transform10), _p(cam0), _p(cam1), _p(depth_map0), _p(variance_map0),
                                   1, H, W, float(default_depth), float(default_variance),
                                   float(uncertaintity_bias), _p(d1), _p(v1)))
    return d1, v1


def increment_age(age_map0, camera_params0, camera_params1, transform10, depth_map0):
    """-> uint64 [H,W] (src/py/semi_dense.rs:35-51; note cameras come BEFORE the transform).
    One host-pointer C-ABI call (tdk_increment_age_host)."""
    age_map0 = _need(age_map0, "age_map0", np.uint64, 2)
    transform10 = _need(transform10, "transform10", np.float64, 2)
    depth_map0 = _need(depth_map0, "depth_map0", np.float64, 2)
    if transform10.shape != (4, 4) or depth_map0.shape != age_map0.shape:
        raise ValueError("bad shapes")
    device._require_cuda()
    H, W = age_map0.shape
    cam0, cam1 = camera_vector(camera_params0), camera_vector(camera_params1)
    out = np.empty((H, W), dtype=np.uint64)
    _lib.check(_lib.increment_age_host(_p(age_map0), _p(cam0), _p(cam1), _p(transform10), _p(depth_map0),
                                       1, H, W, _p(out)))
    return out


def fusion(mu1, mu2, var1, var2):
    """Elementwise Gaussian fusion -> (mu, var) (src/semi_dense/fusion.rs:13-42; argument order of
    `fusion_arrays` and of examples/semi_dense_depth.py:41-42).  tdk_fusion_host."""
    arrs = [_need(a, n, np.float64, 2) for a, n in
            ((mu1, "mu1"), (mu2, "mu2"), (var1, "var1"), (var2, "var2"))]
    if any(a.shape != arrs[0].shape for a in arrs):
        raise ValueError("shapes differ")
    device._require_cuda()
    mu, var = np.empty(arrs[0].shape), np.empty(arrs[0].shape)
    if mu.size:
        _lib.check(_lib.fusion_host(*[_p(a) for a in arrs], None, mu.size, 0, _p(mu), _p(var)))
    return mu, var


def regularize(depth_map, variance_map, flag_map):
    """3x3 validity-weighted mean of inverse depth -> depth map (the signature of the reference's
    commented-out binding, src/py/semi_dense.rs:220-233).  tdk_regularize_host."""
    depth_map = _need(depth_map, "depth_map", np.float64, 2)
    variance_map = _need(variance_map, "variance_map", np.float64, 2)
    flag_map = _need(flag_map, "flag_map", np.int64, 2)
    if depth_map.shape != variance_map.shape or depth_map.shape != flag_map.shape:
        raise ValueError("shapes differ")
    device._require_cuda()
    H, W = depth_map.shape
    out = np.empty((H, W))
    _lib.check(_lib.regularize_host(_p(depth_map), _p(variance_map), _p(flag_map), 1, H, W, _p(out)))
    return out
