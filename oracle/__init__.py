"""ctypes front-end of the parity oracle (``oracle/tdk_oracle.c``).

TEST INFRASTRUCTURE ONLY: imported by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.  Nothing under
``tadataka_b200/`` imports this package.

Each wrapper mirrors one reference function; see the C file for the file:line citations.
Parity status: pinned (tests/test_oracle_golden.py replays the reference's own known-answer
vectors; tests/golden/ holds the NewTsukuba flag cases).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libtdk_oracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libref_bilinear.so")

c_double_p = C.POINTER(C.c_double)
c_i64_p = C.POINTER(C.c_int64)
c_u64_p = C.POINTER(C.c_uint64)


def build(ref=True):
    """Compile the C restatement (and, when /root/reference is present, oracle/_ref)."""
    subprocess.check_call(["make", "-C", _HERE, "-s", "all"])
    if ref:
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"])


def _load():
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(
            os.path.join(_HERE, "tdk_oracle.c")):
        build(ref=False)
    return C.CDLL(_SO)


_lib = _load()


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _dp(a):
    return a.ctypes.data_as(c_double_p)


def _ip(a):
    return a.ctypes.data_as(c_i64_p)


def _up(a):
    return a.ctypes.data_as(c_u64_p)


def _sig(name, restype, *argtypes):
    f = getattr(_lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


D, I64, INT = C.c_double, C.c_int64, C.c_int
DP, IP, UP = c_double_p, c_i64_p, c_u64_p

_safe_invert = _sig("ora_safe_invert", D, D)
_clamp = _sig("ora_clamp", D, D, D, D)
_norm = _sig("ora_norm", D, INT, DP)
_normalize = _sig("ora_normalize", None, INT, DP, DP)
_project = _sig("ora_project", None, DP, DP)
_inv_project = _sig("ora_inv_project", None, DP, D, DP)
_transform = _sig("ora_transform", None, DP, DP, DP)
_warp = _sig("ora_warp", None, DP, DP, D, DP, DP)
_cam_normalize = _sig("ora_cam_normalize", None, DP, DP, DP)
_cam_unnormalize = _sig("ora_cam_unnormalize", None, DP, DP, DP)
_perspective_warp = _sig("ora_perspective_warp", None, DP, DP, DP, DP, D, DP, DP)
_mat4_inverse = _sig("ora_mat4_inverse", INT, DP, DP)
_transform_rk = _sig("ora_transform_rk", INT, DP, DP, DP)
_inv_transform = _sig("ora_inv_transform", None, DP, DP)
_calc_key_epipole = _sig("ora_calc_key_epipole", None, DP, DP, DP)
_key_coordinates = _sig("ora_key_coordinates", None, DP, DP, D, DP)
_ref_coordinates = _sig("ora_ref_coordinates", I64, DP, DP, D, DP, I64)
_is_in_range = _sig("ora_is_in_range", INT, D, D, I64, I64)
_interpolate_n = _sig("ora_interpolate_n", None, DP, I64, DP, I64, DP)
_sobel = _sig("ora_sobel", None, DP, I64, I64, DP, DP)
_gradient1d = _sig("ora_gradient1d", None, DP, INT, DP)
_check_args = _sig("ora_check_args", INT, D, D, D, D)
_hypothesis_range = _sig("ora_hypothesis_range", None, D, D, D, D, DP, DP)
_depth_search_range = _sig("ora_depth_search_range", None, D, D, DP, DP)
_calc_ref_depth = _sig("ora_calc_ref_depth", D, DP, DP, D)
_step_ratio = _sig("ora_step_ratio", INT, DP, DP, D, DP)
_calc_key_direction = _sig("ora_calc_key_direction", None, DP, DP, DP, DP)
_calc_ref_ends = _sig("ora_calc_ref_ends", None, DP, DP, D, D, DP, DP)
_check_us_ref = _sig("ora_check_us_ref", INT, I64, DP, DP, I64, I64, I64)
_search = _sig("ora_search", I64, DP, I64, DP, INT)
_calc_depth0 = _sig("ora_calc_depth0", D, DP, DP, DP)
_calc_variance = _sig("ora_calc_variance", D, D, D, D, D, D)
_photo_var = _sig("ora_photo_var", D, D)
_geo_var_ = _sig("ora_geo_var_", D, DP, DP)
_geo_var = _sig("ora_geo_var", D, DP, DP, DP)
_alpha_ = _sig("ora_alpha_", D, DP, D, D, DP, DP, D, D)
_ref_epipolar_direction = _sig("ora_ref_epipolar_direction", None, DP, DP, D, D, DP)
_calc_alpha_ = _sig("ora_calc_alpha_", D, DP, DP, DP, D)
_calc_alpha = _sig("ora_calc_alpha", D, DP, DP, D, D, D)
_estimate_debug = _sig("ora_estimate_debug", INT, IP, D, D, DP, DP, DP, DP, DP, DP, I64, I64, DP,
                       DP, DP, IP)
_update_depth_batch = _sig("ora_update_depth_batch", INT, DP, DP, DP, DP, DP, DP, I64, I64, UP,
                           DP, DP, DP, I64, I64, DP, DP, IP, IP, IP, INT)
_fusion_arrays = _sig("ora_fusion_arrays", None, DP, DP, DP, DP, I64, DP, DP)
_are_statically_same = _sig("ora_are_statically_same", INT, D, D, D, D)
_propagate_variance = _sig("ora_propagate_variance", D, D, D, D, D)
_handle_collision = _sig("ora_handle_collision", None, D, D, D, D, DP, DP)
_propagate = _sig("ora_propagate", None, DP, DP, DP, DP, DP, I64, I64, D, D, D, DP, DP)
_increment_age = _sig("ora_increment_age", None, UP, DP, DP, DP, DP, I64, I64, UP)
_regularize = _sig("ora_regularize", None, DP, DP, IP, I64, I64, DP)
_warp_vecs = _sig("ora_warp_vecs", None, DP, DP, DP, I64, DP, DP)
_transform_vecs = _sig("ora_transform_vecs", None, DP, DP, I64, DP)
_project_vecs = _sig("ora_project_vecs", None, DP, I64, DP)
_inv_project_vecs = _sig("ora_inv_project_vecs", None, DP, DP, I64, DP)


# ------------------------------------------------------------------ scalar / small helpers
def safe_invert(v):
    return _safe_invert(v)


def clamp(v, lo, hi):
    return _clamp(v, lo, hi)


def norm(v):
    v = _f64(v)
    return _norm(v.size, _dp(v))


def normalize(v):
    v = _f64(v)
    out = np.empty_like(v)
    _normalize(v.size, _dp(v), _dp(out))
    return out


def project(p):
    p = _f64(p)
    out = np.empty(2)
    _project(_dp(p), _dp(out))
    return out


def inv_project(x, depth):
    x = _f64(x)
    out = np.empty(3)
    _inv_project(_dp(x), depth, _dp(out))
    return out


def transform(T, p):
    T, p = _f64(T), _f64(p)
    out = np.empty(3)
    _transform(_dp(T), _dp(p), _dp(out))
    return out


def warp(T, x0, depth0):
    T, x0 = _f64(T), _f64(x0)
    x1 = np.empty(2)
    d1 = C.c_double()
    _warp(_dp(T), _dp(x0), depth0, _dp(x1), C.byref(d1))
    return x1, d1.value


def cam(focal_length, offset):
    return np.array([focal_length[0], focal_length[1], offset[0], offset[1]], dtype=np.float64)


def cam_normalize(camera, u):
    camera, u = _f64(camera), _f64(u)
    out = np.empty(2)
    _cam_normalize(_dp(camera), _dp(u), _dp(out))
    return out


def cam_unnormalize(camera, x):
    camera, x = _f64(camera), _f64(x)
    out = np.empty(2)
    _cam_unnormalize(_dp(camera), _dp(x), _dp(out))
    return out


def perspective_warp(T10, cam0, cam1, u0, depth0):
    T10, cam0, cam1, u0 = _f64(T10), _f64(cam0), _f64(cam1), _f64(u0)
    u1 = np.empty(2)
    d1 = C.c_double()
    _perspective_warp(_dp(T10), _dp(cam0), _dp(cam1), _dp(u0), depth0, _dp(u1), C.byref(d1))
    return u1, d1.value


def mat4_inverse(A):
    A = _f64(A)
    out = np.empty((4, 4))
    info = _mat4_inverse(_dp(A), _dp(out))
    return out, info


def transform_rk(T_wk, T_wr):
    T_wk, T_wr = _f64(T_wk), _f64(T_wr)
    out = np.empty((4, 4))
    _transform_rk(_dp(T_wk), _dp(T_wr), _dp(out))
    return out


def inv_transform(T):
    T = _f64(T)
    out = np.empty((4, 4))
    _inv_transform(_dp(T), _dp(out))
    return out


def calc_key_epipole(T_wk, T_wr):
    T_wk, T_wr = _f64(T_wk), _f64(T_wr)
    out = np.empty(2)
    _calc_key_epipole(_dp(T_wk), _dp(T_wr), _dp(out))
    return out


def key_coordinates(direction, x_key, step_size):
    direction, x_key = _f64(direction), _f64(x_key)
    out = np.empty((5, 2))
    _key_coordinates(_dp(direction), _dp(x_key), step_size, _dp(out))
    return out


def ref_coordinates(x_min, direction, step_size, cap=1 << 20):
    x_min, direction = _f64(x_min), _f64(direction)
    out = np.empty((cap, 2))
    n = _ref_coordinates(_dp(x_min), _dp(direction), step_size, _dp(out), cap)
    return out[:min(n, cap)].copy()


def is_in_range(x, y, shape):
    return bool(_is_in_range(x, y, shape[0], shape[1]))


def interpolate(image, coords):
    image = _f64(image)
    coords = _f64(coords).reshape(-1, 2)
    out = np.empty(coords.shape[0])
    _interpolate_n(_dp(image), image.shape[1], _dp(coords), coords.shape[0], _dp(out))
    return out


def sobel(image):
    image = _f64(image)
    H, W = image.shape
    gx, gy = np.empty((H, W)), np.empty((H, W))
    _sobel(_dp(image), H, W, _dp(gx), _dp(gy))
    return gx, gy


def gradient1d(x):
    x = _f64(x)
    out = np.empty(x.size - 1)
    _gradient1d(_dp(x), x.size, _dp(out))
    return out


def check_args(inv_depth, variance, inv_depth_range):
    return _check_args(inv_depth, variance, inv_depth_range[0], inv_depth_range[1])


def hypothesis_range(inv_depth, variance, inv_depth_range):
    lo, hi = C.c_double(), C.c_double()
    _hypothesis_range(inv_depth, variance, inv_depth_range[0], inv_depth_range[1],
                      C.byref(lo), C.byref(hi))
    return lo.value, hi.value


def depth_search_range(min_inv, max_inv):
    a, b = C.c_double(), C.c_double()
    _depth_search_range(min_inv, max_inv, C.byref(a), C.byref(b))
    return a.value, b.value


def calc_ref_depth(T_rk, x_key, depth_key):
    T_rk, x_key = _f64(T_rk), _f64(x_key)
    return _calc_ref_depth(_dp(T_rk), _dp(x_key), depth_key)


def step_ratio(T_rk, x_key, key_inv_depth):
    T_rk, x_key = _f64(T_rk), _f64(x_key)
    out = C.c_double()
    flag = _step_ratio(_dp(T_rk), _dp(x_key), key_inv_depth, C.byref(out))
    return flag, out.value


def calc_key_direction(x_key, e_key, ref_dir):
    x_key, e_key, ref_dir = _f64(x_key), _f64(e_key), _f64(ref_dir)
    out = np.empty(2)
    _calc_key_direction(_dp(x_key), _dp(e_key), _dp(ref_dir), _dp(out))
    return out


def calc_ref_ends(T_rk, x_key, depth_range):
    T_rk, x_key = _f64(T_rk), _f64(x_key)
    a, b = np.empty(2), np.empty(2)
    _calc_ref_ends(_dp(T_rk), _dp(x_key), depth_range[0], depth_range[1], _dp(a), _dp(b))
    return a, b


def check_us_ref(us_ref, us_key_size, shape):
    us_ref = _f64(us_ref).reshape(-1, 2)
    n = us_ref.shape[0]
    first = us_ref[0].copy() if n else np.zeros(2)
    last = us_ref[-1].copy() if n else np.zeros(2)
    return _check_us_ref(n, _dp(first), _dp(last), us_key_size, shape[0], shape[1])


def search(sequence, kernel):
    sequence, kernel = _f64(sequence), _f64(kernel)
    return _search(_dp(sequence), sequence.size, _dp(kernel), kernel.size)


def calc_depth0(T10, x0, x1):
    T10, x0, x1 = _f64(T10), _f64(x0), _f64(x1)
    return _calc_depth0(_dp(T10), _dp(x0), _dp(x1))


def calc_variance(alpha, geo_var, photo_var, geo, photo):
    return _calc_variance(alpha, geo_var, photo_var, geo, photo)


def photo_var(gradient):
    return _photo_var(gradient)


def geo_var_(direction, image_grad):
    direction, image_grad = _f64(direction), _f64(image_grad)
    return _geo_var_(_dp(direction), _dp(image_grad))


def geo_var(x_key, t_rk, image_grad):
    x_key, t_rk, image_grad = _f64(x_key), _f64(t_rk), _f64(image_grad)
    return _geo_var(_dp(x_key), _dp(t_rk), _dp(image_grad))


def alpha_(x_key, x_ref_i, direction_i, ri, rz, ti, tz):
    x_key, ri, rz = _f64(x_key), _f64(ri), _f64(rz)
    return _alpha_(_dp(x_key), x_ref_i, direction_i, _dp(ri), _dp(rz), ti, tz)


def ref_epipolar_direction(T_rk, x_key, depth_range):
    T_rk, x_key = _f64(T_rk), _f64(x_key)
    out = np.empty(2)
    _ref_epipolar_direction(_dp(T_rk), _dp(x_key), depth_range[0], depth_range[1], _dp(out))
    return out


def calc_alpha_(T_rk, x_key, direction, prior_depth):
    T_rk, x_key, direction = _f64(T_rk), _f64(x_key), _f64(direction)
    return _calc_alpha_(_dp(T_rk), _dp(x_key), _dp(direction), prior_depth)


def calc_alpha(T_rk, x_key, depth_range, prior_depth):
    T_rk, x_key = _f64(T_rk), _f64(x_key)
    return _calc_alpha(_dp(T_rk), _dp(x_key), depth_range[0], depth_range[1], prior_depth)


# ------------------------------------------------------------------ the path itself
def make_params(min_depth, max_depth, geo_coeff, photo_coeff, ref_step_size, min_gradient):
    """src/py/semi_dense.rs:91-109 -- note the range is stored inverted."""
    return np.array([safe_invert(max_depth), safe_invert(min_depth), geo_coeff, photo_coeff,
                     ref_step_size, min_gradient], dtype=np.float64)


def estimate_debug(u_key, prior_depth, prior_variance, key_img, key_cam, T_wk, ref_img, ref_cam,
                   T_wr, params, with_debug=False):
    """src/py/semi_dense.rs:126-155 -> (depth, variance, flag[, (argmin, n)])."""
    u = np.ascontiguousarray(u_key, dtype=np.int64)
    key_img, ref_img = _f64(key_img), _f64(ref_img)
    key_cam, ref_cam, T_wk, T_wr, params = map(_f64, (key_cam, ref_cam, T_wk, T_wr, params))
    H, W = key_img.shape
    d, v = C.c_double(), C.c_double()
    dbg = np.full(2, -1, dtype=np.int64)
    flag = _estimate_debug(_ip(u), prior_depth, prior_variance, _dp(key_img), _dp(key_cam),
                           _dp(T_wk), _dp(ref_img), _dp(ref_cam), _dp(T_wr), H, W, _dp(params),
                           C.byref(d), C.byref(v), _ip(dbg))
    if with_debug:
        return d.value, v.value, flag, (int(dbg[0]), int(dbg[1]))
    return d.value, v.value, flag


def update_depth_batch(key_imgs, key_cams, T_wks, ref_imgs, ref_cams, T_wrs, age_maps,
                       prior_depth, prior_variance, params, threads=1, debug=False):
    """P keyframes x R refs.  Shapes: key_imgs [P,H,W], key_cams [P,4], T_wks [P,4,4],
    ref_imgs [P,R,H,W], ref_cams [P,R,4], T_wrs [P,R,4,4], maps [P,H,W].
    Returns (depth, variance, flag[, argmin, n]); raises ValueError if an age exceeds R."""
    key_imgs, ref_imgs = _f64(key_imgs), _f64(ref_imgs)
    P, H, W = key_imgs.shape
    R = ref_imgs.shape[1]
    key_cams, T_wks, ref_cams, T_wrs = map(_f64, (key_cams, T_wks, ref_cams, T_wrs))
    age_maps = np.ascontiguousarray(age_maps, dtype=np.uint64)
    prior_depth, prior_variance, params = map(_f64, (prior_depth, prior_variance, params))
    assert ref_imgs.shape == (P, R, H, W) and age_maps.shape == (P, H, W)
    assert prior_depth.shape == (P, H, W) and prior_variance.shape == (P, H, W)
    out_d, out_v = np.empty((P, H, W)), np.empty((P, H, W))
    out_f = np.empty((P, H, W), dtype=np.int64)
    if debug:
        am, nn = np.empty((P, H, W), dtype=np.int64), np.empty((P, H, W), dtype=np.int64)
        amp, nnp = _ip(am), _ip(nn)
    else:
        amp = nnp = None
    rc = _update_depth_batch(_dp(key_imgs), _dp(key_cams), _dp(T_wks), _dp(ref_imgs),
                             _dp(ref_cams), _dp(T_wrs), P, R, _up(age_maps), _dp(prior_depth),
                             _dp(prior_variance), _dp(params), H, W, _dp(out_d), _dp(out_v),
                             _ip(out_f), amp, nnp, int(threads))
    if rc != 0:
        raise ValueError("Age exceeds the refframe size")
    if debug:
        return out_d, out_v, out_f, am, nn
    return out_d, out_v, out_f


def update_depth(key_img, key_cam, T_wk, ref_imgs, ref_cams, T_wrs, age_map, prior_depth,
                 prior_variance, params, debug=False):
    """One keyframe (src/semi_dense/semi_dense.rs:160-234)."""
    r = update_depth_batch(_f64(key_img)[None], _f64(key_cam)[None], _f64(T_wk)[None],
                           _f64(ref_imgs)[None], _f64(ref_cams)[None], _f64(T_wrs)[None],
                           np.asarray(age_map)[None], _f64(prior_depth)[None],
                           _f64(prior_variance)[None], params, 1, debug)
    return tuple(a[0] for a in r)


def fusion_arrays(mu1, mu2, var1, var2):
    mu1, mu2, var1, var2 = map(_f64, (mu1, mu2, var1, var2))
    mu, var = np.empty_like(mu1), np.empty_like(mu1)
    _fusion_arrays(_dp(mu1), _dp(mu2), _dp(var1), _dp(var2), mu1.size, _dp(mu), _dp(var))
    return mu, var


def are_statically_same(id1, id2, v1, v2):
    return bool(_are_statically_same(id1, id2, v1, v2))


def propagate_variance(depth0, depth1, variance0, bias):
    return _propagate_variance(depth0, depth1, variance0, bias)


def handle_collision(depth_a, depth_b, var_a, var_b):
    d, v = C.c_double(), C.c_double()
    _handle_collision(depth_a, depth_b, var_a, var_b, C.byref(d), C.byref(v))
    return d.value, v.value


def propagate(T10, cam0, cam1, depth_map0, variance_map0, default_depth, default_variance, bias):
    T10, cam0, cam1 = map(_f64, (T10, cam0, cam1))
    depth_map0, variance_map0 = _f64(depth_map0), _f64(variance_map0)
    H, W = depth_map0.shape
    d1, v1 = np.empty((H, W)), np.empty((H, W))
    _propagate(_dp(T10), _dp(cam0), _dp(cam1), _dp(depth_map0), _dp(variance_map0), H, W,
               default_depth, default_variance, bias, _dp(d1), _dp(v1))
    return d1, v1


def increment_age(age_map0, cam0, cam1, T10, depth_map0):
    age_map0 = np.ascontiguousarray(age_map0, dtype=np.uint64)
    T10, cam0, cam1, depth_map0 = map(_f64, (T10, cam0, cam1, depth_map0))
    H, W = age_map0.shape
    out = np.empty((H, W), dtype=np.uint64)
    _increment_age(_up(age_map0), _dp(cam0), _dp(cam1), _dp(T10), _dp(depth_map0), H, W, _up(out))
    return out


def regularize(depth_map, variance_map, flag_map):
    depth_map, variance_map = _f64(depth_map), _f64(variance_map)
    flag_map = np.ascontiguousarray(flag_map, dtype=np.int64)
    H, W = depth_map.shape
    out = np.empty((H, W))
    _regularize(_dp(depth_map), _dp(variance_map), _ip(flag_map), H, W, _dp(out))
    return out


def warp_vecs(T10, xs0, depths0):
    T10, xs0, depths0 = _f64(T10), _f64(xs0), _f64(depths0)
    n = depths0.size
    xs1, d1 = np.empty((n, 2)), np.empty(n)
    _warp_vecs(_dp(T10), _dp(xs0), _dp(depths0), n, _dp(xs1), _dp(d1))
    return xs1, d1


def transform_vecs(T10, ps0):
    T10, ps0 = _f64(T10), _f64(ps0)
    out = np.empty_like(ps0)
    _transform_vecs(_dp(T10), _dp(ps0), ps0.shape[0], _dp(out))
    return out


def project_vecs(ps):
    ps = _f64(ps)
    out = np.empty((ps.shape[0], 2))
    _project_vecs(_dp(ps), ps.shape[0], _dp(out))
    return out


def inv_project_vecs(xs, depths):
    xs, depths = _f64(xs), _f64(depths)
    out = np.empty((xs.shape[0], 3))
    _inv_project_vecs(_dp(xs), _dp(depths), xs.shape[0], _dp(out))
    return out


def ref_bilinear(image, coords):
    """The reference's own legacy C++ bilinear sampler (tadataka/interpolation/_bilinear.cpp),
    compiled by `make ref` into oracle/_ref/.  Returns None when that build is absent."""
    if not os.path.exists(_REF_SO):
        return None
    lib = C.CDLL(_REF_SO)
    f = getattr(lib, "_Z14_interpolationPKdiS0_iPd")
    f.restype = None
    f.argtypes = [DP, C.c_int, DP, C.c_int, DP]
    image = _f64(image)
    coords = _f64(coords).reshape(-1, 2)
    out = np.empty(coords.shape[0])
    f(_dp(image), image.shape[1], _dp(coords), coords.shape[0], _dp(out))
    return out
