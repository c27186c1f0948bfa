"""numpy-in / numpy-out forms of the geometry and sampling helpers the reference exposes as
`rust_bindings.{warp,interpolation,projection,transform,triangulation,homogeneous}`
(src/py/warp.rs, interpolation.rs, projection.rs, transform.rs, triangulation.rs, homogeneous.rs).
The arithmetic runs in libtdk_b200.so on the current CUDA device; the homogeneous helpers only
move data (append / drop a column of ones) and stay on the host.
"""
import numpy as np
import torch

from . import device


def _need(a, name, ndim):
    if not isinstance(a, np.ndarray) or a.dtype != np.float64 or a.ndim != ndim:
        raise TypeError(f"{name} must be a {ndim}-D float64 numpy array")
    return np.ascontiguousarray(a)


def _cuda(a):
    device._require_cuda()
    return torch.from_numpy(a).cuda()


# ---- rust_bindings.warp (src/py/warp.rs:23-75)
def warp_vecs(transform10, xs, depths):
    T, xs, depths = _need(transform10, "transform10", 2), _need(xs, "xs", 2), _need(depths, "depths", 1)
    xs1, d1 = device.warp_vecs(_cuda(T), _cuda(xs), _cuda(depths))
    return xs1.cpu().numpy(), d1.cpu().numpy()


def warp_vec(transform10, x0, depth0):
    x0 = _need(x0, "x0", 1)
    xs1, d1 = warp_vecs(transform10, x0[None, :], np.array([depth0], dtype=np.float64))
    return xs1[0], float(d1[0])


# ---- rust_bindings.interpolation (src/py/interpolation.rs:6-22)
def interpolation(image, coordinates):
    image, coordinates = _need(image, "image", 2), _need(coordinates, "coordinates", 2)
    return device.interpolate(_cuda(image), _cuda(coordinates)).cpu().numpy()


# ---- rust_bindings.projection (src/py/projection.rs:7-51)
def project_vecs(xs):
    return device.project_vecs(_cuda(_need(xs, "xs", 2))).cpu().numpy()


def project_vec(x):
    return project_vecs(_need(x, "x", 1)[None, :])[0]


def inv_project_vecs(xs, depths):
    return device.inv_project_vecs(_cuda(_need(xs, "xs", 2)), _cuda(_need(depths, "depths", 1))).cpu().numpy()


def inv_project_vec(x, depth):
    return inv_project_vecs(_need(x, "x", 1)[None, :], np.array([depth], dtype=np.float64))[0]


# ---- rust_bindings.transform (src/py/transform.rs:6-22)
def transform(transform10, points0):
    T, ps = _need(transform10, "transform10", 2), _need(points0, "points0", 2)
    return device.transform(_cuda(T), _cuda(ps)).cpu().numpy()


# ---- rust_bindings.triangulation (src/py/triangulation.rs:7-26)
def calc_depth0(transform_10, x0, x1):
    T, x0, x1 = _need(transform_10, "transform_10", 2), _need(x0, "x0", 1), _need(x1, "x1", 1)
    return float(device.calc_depth0(_cuda(T), _cuda(x0[None, :]), _cuda(x1[None, :])).cpu().numpy()[0])


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
