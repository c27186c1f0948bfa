"""numpy-in / numpy-out forms of the geometry and sampling helpers the reference exposes as
`rust_bindings.{warp,interpolation,projection,transform,triangulation,homogeneous}`
(src/py/warp.rs, interpolation.rs, projection.rs, transform.rs, triangulation.rs, homogeneous.rs).
The arithmetic runs in libtdk_b200.so on the current CUDA device; the homogeneous helpers only
move data (append / drop a column of ones) and stay on the host.
"""
import ctypes as C

import numpy as np

from . import _lib, device


def _need(a, name, ndim):
    if not isinstance(a, np.ndarray) or a.dtype != np.float64 or a.ndim != ndim:
        raise TypeError(f"{name} must be a {ndim}-D float64 numpy array")
    return np.ascontiguousarray(a)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ---- rust_bindings.warp (src/py/warp.rs:23-75)
def warp_vecs(transform10, xs, depths):
    T, xs, depths = _need(transform10, "transform10", 2), _need(xs, "xs", 2), _need(depths, "depths", 1)
    if T.shape != (4, 4) or xs.shape != (depths.shape[0], 2):
        raise ValueError("bad shapes")
    device._require_cuda()
    xs1, d1 = np.empty_like(xs), np.empty_like(depths)
    _lib.check(_lib.warp_vecs_host(_p(T), _p(xs), _p(depths), depths.shape[0], _p(xs1), _p(d1)))
    return xs1, d1


def warp_vec(transform10, x0, depth0):
    x0 = _need(x0, "x0", 1)
    xs1, d1 = warp_vecs(transform10, x0[None, :], np.array([depth0], dtype=np.float64))
    return xs1[0], float(d1[0])


# ---- rust_bindings.interpolation (src/py/interpolation.rs:6-22)
def interpolation(image, coordinates):
    image, coordinates = _need(image, "image", 2), _need(coordinates, "coordinates", 2)
    if coordinates.shape[1] != 2:
        raise ValueError("coordinates must be [N,2]")
    device._require_cuda()
    out = np.empty(coordinates.shape[0])
    rc = _lib.interpolate_host(_p(image), image.shape[0], image.shape[1], _p(coordinates), coordinates.shape[0], _p(out))
    if rc == _lib.ERR_RANGE:
        raise ValueError("Coordinates {} out of image range".format("[...]"))
    _lib.check(rc)
    return out


# ---- rust_bindings.projection (src/py/projection.rs:7-51)
def project_vecs(xs):
    ps = _need(xs, "xs", 2)
    if ps.shape[1] != 3:
        raise ValueError("points must be [N,3]")
    device._require_cuda()
    out = np.empty((ps.shape[0], 2))
    _lib.check(_lib.project_vecs_host(_p(ps), ps.shape[0], _p(out)))
    return out


def project_vec(x):
    return project_vecs(_need(x, "x", 1)[None, :])[0]


def inv_project_vecs(xs, depths):
    xs, depths = _need(xs, "xs", 2), _need(depths, "depths", 1)
    if xs.shape != (depths.shape[0], 2):
        raise ValueError("bad shapes")
    device._require_cuda()
    out = np.empty((xs.shape[0], 3))
    _lib.check(_lib.inv_project_vecs_host(_p(xs), _p(depths), xs.shape[0], _p(out)))
    return out


def inv_project_vec(x, depth):
    return inv_project_vecs(_need(x, "x", 1)[None, :], np.array([depth], dtype=np.float64))[0]


# ---- rust_bindings.transform (src/py/transform.rs:6-22)
def transform(transform10, points0):
    T, ps = _need(transform10, "transform10", 2), _need(points0, "points0", 2)
    if T.shape != (4, 4) or ps.shape[1] != 3:
        raise ValueError("bad shapes")
    device._require_cuda()
    out = np.empty_like(ps)
    _lib.check(_lib.transform_host(_p(T), _p(ps), ps.shape[0], _p(out)))
    return out


# ---- rust_bindings.triangulation (src/py/triangulation.rs:7-26)
def calc_depth0(transform_10, x0, x1):
    T, x0, x1 = _need(transform_10, "transform_10", 2), _need(x0, "x0", 1), _need(x1, "x1", 1)
    device._require_cuda()
    out = np.empty(1)
    _lib.check(_lib.calc_depth0_host(_p(T), _p(x0), _p(x1), 1, _p(out)))
    return float(out[0])


# ---- rust_bindings.homogeneous (src/py/homogeneous.rs:6-44)
def to_homogeneous_vec(x):
    return np.concatenate([_need(x, "x", 1), np.ones(1)])


def to_homogeneous_vecs(xs):
    xs = _need(xs, "xs", 2)
    return np.concatenate([xs, np.ones((xs.shape[0], 1))], axis=1)


def from_homogeneous_vec(x):
    return _need(x, "x", 1)[:-1].copy()


def from_homogeneous_vecs(xs):
    return _need(xs, "xs", 2)[:, :-1].copy()
