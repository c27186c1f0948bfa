"""`rust_bindings.camera` (reference: src/py/camera.rs:6-41, src/camera.rs:4-15)."""
import numpy as np


class CameraParameters:
    """Pinhole intrinsics.  `CameraParameters((fx, fy), (ox, oy))`; the getters return fresh
    float64 arrays of length 2, as the pyo3 class does."""

    def __init__(self, focal_length, offset):
        fx, fy = (float(v) for v in focal_length)
        ox, oy = (float(v) for v in offset)
        self._v = np.array([fx, fy, ox, oy], dtype=np.float64)

    @property
    def focal_length(self):
        return self._v[0:2].copy()

    @property
    def offset(self):
        return self._v[2:4].copy()

    def __repr__(self):
        return "CameraParameters(focal_length={}, offset={})".format(
            tuple(self._v[0:2]), tuple(self._v[2:4]))


def camera_vector(camera_params):
    """{fx, fy, ox, oy} from ANY object with float64 `.focal_length` / `.offset` arrays
    (duck-typed exactly like src/py/semi_dense.rs:19-33, including the dtype strictness)."""
    f = getattr(camera_params, "focal_length")
    o = getattr(camera_params, "offset")
    for name, a in (("focal_length", f), ("offset", o)):
        if not isinstance(a, np.ndarray) or a.dtype != np.float64 or a.ndim != 1:
            raise TypeError(f"camera_params.{name} must be a 1-D float64 numpy array")
        if a.shape[0] < 2:
            raise IndexError(f"camera_params.{name} needs two elements")
    return np.array([f[0], f[1], o[0], o[1]], dtype=np.float64)
