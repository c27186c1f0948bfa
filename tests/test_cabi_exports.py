"""The C-ABI library loads and exports every symbol include/tdk.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "tdk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tdk_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from tadataka_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tdk.h but not exported"
    assert sorted(_lib.EXPORTS) == names


def test_version_and_error_text_without_a_gpu():
    from tadataka_b200 import _lib
    assert b"sm_100a" in _lib.version()
    # argument validation happens before any CUDA call
    assert _lib.set_kernel_plan(7) == _lib.ERR_INVALID
    assert b"not in {-1,0,1,2}" in _lib.last_error()
    assert _lib.update_depth_workspace_bytes(0, 1, 480, 640) == 0
    assert _lib.update_depth_workspace_bytes(4, 8, 480, 640) >= 4 * (4 + 133) * 480 * 640
    assert _lib.propagate_workspace_bytes(1, 480, 640) >= 28 * 480 * 640


def test_reference_module_surface():
    """The eight rust_bindings modules and the names the reference's Python imports from them
    (src/py/*.rs #[pymodule] blocks)."""
    import importlib
    surface = {
        "semi_dense": ["Frame", "Params", "estimate_debug_", "increment_age", "update_depth",
                       "propagate"],
        "camera": ["CameraParameters"],
        "warp": ["warp_vec", "warp_vecs"],
        "interpolation": ["interpolation"],
        "projection": ["project_vec", "project_vecs", "inv_project_vec", "inv_project_vecs"],
        "transform": ["transform"],
        "triangulation": ["calc_depth0"],
        "homogeneous": ["to_homogeneous_vec", "to_homogeneous_vecs", "from_homogeneous_vec",
                        "from_homogeneous_vecs"],
    }
    for mod, names in surface.items():
        m = importlib.import_module("rust_bindings." + mod)
        for n in names:
            assert hasattr(m, n), f"rust_bindings.{mod}.{n} missing"
    from tadataka.vo.semi_dense.flag import ResultFlag
    assert [int(f) for f in ResultFlag] == list(range(0, -10, -1))


def test_frame_and_params_constructors():
    # tests/vo/semi_dense/test_semi_dense.py:16-31
    import numpy as np
    import pytest
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import Frame, Params
    cp = CameraParameters((10., 10.), (20., 10.))
    frame = Frame(cp, np.zeros((40, 20)), np.zeros((4, 4)))
    assert frame.image.shape == (40, 20) and frame.transform_wf.shape == (4, 4)
    assert frame.camera_params.focal_length.tolist() == [10., 10.]
    assert frame.camera_params.offset.dtype == np.float64
    Params(min_depth=0.1, max_depth=10.0, geo_coeff=0.4, photo_coeff=0.5, ref_step_size=0.02,
           min_gradient=0.001)
    Params(0.1, 10.0, 0.4, 0.5, 0.02, 0.001)
    with pytest.raises(TypeError):     # exact-dtype extraction, like pyo3-numpy
        Frame(cp, np.zeros((40, 20), dtype=np.float32), np.zeros((4, 4)))
    # any object with float64 .focal_length / .offset works (duck typing, py/semi_dense.rs:19-33)

    class Duck:
        focal_length = np.array([1., 2.])
        offset = np.array([3., 4.])
    assert Frame(Duck(), np.zeros((4, 4)), np.eye(4)).camera_params.offset.tolist() == [3., 4.]


def test_argument_errors_are_raised_before_any_cuda_call():
    """Exact-dtype and shape checks of the drop-in layer (pyo3-numpy never casts:
    src/py/semi_dense.rs:38-43,158-166); these fire before the GPU is touched."""
    import numpy as np
    import pytest
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import (Frame, Params, estimate_debug_, increment_age, propagate,
                                          update_depth)
    import rust_bindings.warp as rw
    cp = CameraParameters((10., 10.), (8., 6.))
    frame = Frame(cp, np.zeros((12, 16)), np.eye(4))
    params = Params(0.1, 10.0, 0.4, 0.5, 0.02, 0.001)
    age, d, v = np.ones((12, 16), dtype=np.uint64), np.ones((12, 16)), np.ones((12, 16))
    with pytest.raises(TypeError):
        update_depth(frame, [frame], age.astype(np.int64), d, v, params)       # age must be uint64
    with pytest.raises(TypeError):
        update_depth(frame, [frame], age, d.astype(np.float32), v, params)
    with pytest.raises(TypeError):
        update_depth(frame, [object()], age, d, v, params)                      # not a Frame
    with pytest.raises(TypeError):
        update_depth(frame, [frame], age, d, v, (0.1, 10.0, 0.4, 0.5, 0.02, 0.001))
    with pytest.raises(ValueError):
        update_depth(frame, [frame], age[:5], d, v, params)                     # shape mismatch
    with pytest.raises(ValueError):
        update_depth(frame, [Frame(cp, np.zeros((6, 16)), np.eye(4))], age, d, v, params)
    with pytest.raises(TypeError):
        estimate_debug_(np.array([1., 2.]), 1.0, 1.0, frame, frame, params)     # u_key must be int64
    with pytest.raises(TypeError):
        increment_age(age.astype(np.int32), cp, cp, np.eye(4), d)
    with pytest.raises(TypeError):
        propagate(np.eye(4, dtype=np.float32), cp, cp, d, v, 1.0, 1.0, 0.1)
    with pytest.raises(TypeError):
        rw.warp_vecs(np.eye(4), np.zeros((3, 2), dtype=np.float32), np.ones(3))
    # no reference frames: every pixel must be age 0, and comes back unprocessed
    out_d, out_v, out_f = update_depth(frame, [], np.zeros((12, 16), dtype=np.uint64), d, v, params)
    assert (out_f == -9).all() and np.array_equal(out_d, d)
    with pytest.raises(ValueError, match="Age exceeds"):
        update_depth(frame, [], age, d, v, params)


def test_square_root_threshold_is_exact():
    """tdk_sqrt_threshold(m) must be the smallest double T with sqrt(T) >= m, so that the sieve's
    `s < T` is the reference's `sqrt(s) < m` for every s (host arithmetic, no GPU needed)."""
    import math
    import numpy as np
    from tadataka_b200 import _lib
    rng = np.random.default_rng(5)
    ms = list(10.0 ** rng.uniform(-300, 300, 300)) + list(rng.uniform(0.0, 2.0, 300))
    ms += [0.2, 0.05, 1.0, 2.0 ** -537, 2.0 ** 511.5, 5e-324, 1.7976931348623157e308, 1e-162, 1.4e-162]
    for m in ms:
        T = _lib.sqrt_threshold(m)
        assert math.sqrt(T) >= m, (m, T)
        below = np.nextafter(T, 0.0)
        assert T == 0.0 or math.sqrt(below) < m, (m, T)
    assert _lib.sqrt_threshold(0.0) == 0.0 and _lib.sqrt_threshold(-3.0) == 0.0
    assert math.isnan(_lib.sqrt_threshold(float("nan")))
    assert _lib.sqrt_threshold(float("inf")) == float("inf")


def test_update_depth_workspace_sizes():
    from tadataka_b200 import _lib
    H, W, R = 480, 640, 8
    full = [_lib.update_depth_workspace_bytes(P, R, H, W) for P in (1, 2, 64, 65, 300)]
    least = [_lib.update_depth_workspace_bytes_min(P, R, H, W) for P in (1, 2, 64, 65, 300)]
    assert all(a >= b > 0 for a, b in zip(full, least)) and full[0] == least[0]
    assert full == sorted(full) and least == sorted(least)
    # beyond 64 keyframes (one sub-batch) only the per-keyframe constants grow: every per-pixel array
    # belongs to the sub-batch and is reused
    per_kf = (full[4] - full[3]) / (300 - 65)
    assert 0 < per_kf < 4096
    # per sub-batch keyframe: queue 4 + f32 key image 4 + length class 1 + sort position 4 + record 128
    assert 141 * H * W <= (full[2] - full[1]) / 62 < 142 * H * W
