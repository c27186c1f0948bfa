"""Pins the oracle: replays every known-answer vector the reference's own Rust unit tests hold for
the semi-dense path.  Each test names the reference test it replays
(paths relative to /root/reference).  `assert_eq!` in Rust -> exact equality here.
"""
import numpy as np
import pytest
from numpy.testing import assert_array_equal

import oracle as ora

INF = float("inf")


# ---------------------------------------------------------------- src/*.rs helpers
def test_numeric_inv():
    # src/semi_dense/numeric.rs test_inv ; src/numeric.rs
    assert ora.safe_invert(10.0) == 0.1
    assert ora.safe_invert(0.0) == 1.0 / np.finfo(np.float64).eps


def test_clamp():
    # src/cmp.rs test_clamp
    assert ora.clamp(2., 1., 3.) == 2.
    assert ora.clamp(0., 1., 3.) == 1.
    assert ora.clamp(4., 1., 3.) == 3.


def test_normalize():
    # src/vector.rs test_normalize
    assert_array_equal(ora.normalize([4., 3.]), np.array([4., 3.]) / 5.0)
    assert_array_equal(ora.normalize([-1., 2.]), np.array([-1., 2.]) / np.sqrt(5.0))
    assert_array_equal(ora.normalize([0., 0.]), [0., 0.])


def test_projection():
    # src/projection.rs test_projection / test_inv_projection
    pts = np.array([[0., 0., 0.], [1., 4., 2.], [-1., 3., 5.]])
    assert_array_equal(ora.project_vecs(pts), [[0., 0.], [0.5, 2.], [-0.2, 0.6]])
    assert_array_equal(ora.project([0., 0., 0.]), [0., 0.])
    assert_array_equal(ora.project([3., 5., 5.]), [0.6, 1.])
    xs = np.array([[0.5, 2.], [-0.2, 0.6]])
    assert_array_equal(ora.inv_project_vecs(xs, [2., 5.]), [[1., 4., 2.], [-1., 3., 5.]])
    assert_array_equal(ora.inv_project([0.5, 2.], 2.), [1., 4., 2.])


def test_transform():
    # src/transform.rs test_transform_points / test_transform_point / test_inv_transform
    T = np.array([[1, 0, 0, 1], [0, 0, -1, 2], [0, 1, 0, 3], [0, 0, 0, 1]], dtype=float)
    assert_array_equal(ora.transform_vecs(T, [[1., 2., 5.], [4., -2., 3.]]),
                       [[2., -3., 5.], [5., -1., 1.]])
    assert_array_equal(ora.transform(T, [1., 2., 5.]), [2., -3., 5.])
    T = np.array([[1., 0., 0., -1.], [0., 0., 1., -2.], [0., -1., 0., 3.], [0., 0., 0., 1.]])
    assert_array_equal(ora.inv_transform(T) @ T, np.eye(4))


def test_mat4_inverse_against_numpy():
    # third-party arithmetic (ndarray-linalg inv -> LAPACK): restated, checked against numpy's LAPACK
    rng = np.random.default_rng(0)
    for _ in range(20):
        A = rng.normal(size=(4, 4))
        inv, info = ora.mat4_inverse(A)
        assert info == 0
        np.testing.assert_allclose(inv, np.linalg.inv(A), rtol=1e-10, atol=1e-12)
    # needs row interchanges
    A = np.array([[0., 0., 1., 3.], [0., 1., 0., 2.], [-1., 0., 0., 4.], [0., 0., 0., 1.]])
    inv, info = ora.mat4_inverse(A)
    assert info == 0
    assert_array_equal(inv @ A, np.eye(4))
    _, info = ora.mat4_inverse(np.zeros((4, 4)))
    assert info != 0


def test_camera_normalizer():
    # src/camera.rs test_normalizer
    cam = ora.cam((10., 20.), (2., 4.))
    for xn, un in [([1.0, 1.0], [12., 24.]), ([-0.2, -0.2], [0., 0.]), ([0.6, 0.3], [8., 10.])]:
        assert_array_equal(ora.cam_normalize(cam, un), xn)
        assert_array_equal(ora.cam_unnormalize(cam, xn), un)


def test_image_range():
    # src/image_range.rs test_is_in_range / test_all_in_range
    shape = (30, 20)
    assert ora.is_in_range(0., 0., shape)
    for x, y in [(-1., 0.), (0., -1.), (0., 30.), (20., 0.)]:
        assert not ora.is_in_range(x, y, shape)
    pts = [[19., 29.], [19., 0.], [0., 29.], [-1., 29.], [19., -1.], [20., 29.], [19., 30.],
           [20., 30.]]
    assert [ora.is_in_range(x, y, shape) for x, y in pts] == [True] * 3 + [False] * 5
    pts = [[19.00, 29.00], [19.01, 29.00], [19.00, 29.01], [19.01, 29.01], [0., 0.], [0., -0.01],
           [-0.01, 0.], [-0.01, -0.01]]
    assert [ora.is_in_range(x, y, shape) for x, y in pts] == [True, False, False, False, True,
                                                             False, False, False]


IMG = np.array([[0., 1., 5.], [0., 0., 2.], [4., 3., 2.], [5., 6., 1.]])


def _interp_cases():
    i = IMG
    return [
        ([1.3, 2.6], i[2, 1] * (2.0 - 1.3) * (3.0 - 2.6) + i[2, 2] * (1.3 - 1.0) * (3.0 - 2.6)
         + i[3, 1] * (2.0 - 1.3) * (2.6 - 2.0) + i[3, 2] * (1.3 - 1.0) * (2.6 - 2.0)),
        ([0.0, 0.0], i[0, 0]),
        ([0.0, 0.1], i[0, 0] * (1.0 - 0.0) * (1.0 - 0.1) + i[1, 0] * (1.0 - 0.0) * (0.1 - 0.0)),
        ([0.1, 0.0], i[0, 0] * (1.0 - 0.1) * (1.0 - 0.0) + i[0, 1] * (0.1 - 0.0) * (1.0 - 0.0)),
        ([2.0, 2.9], i[2, 2] * (3.0 - 2.0) * (3.0 - 2.9) + i[3, 2] * (3.0 - 2.0) * (2.9 - 2.0)),
        ([1.9, 3.0], i[3, 1] * (2.0 - 1.9) * (4.0 - 3.0) + i[3, 2] * (1.9 - 1.0) * (4.0 - 3.0)),
        ([2.0, 3.0], i[3, 2]),
    ]


def test_interpolation():
    # src/interpolation.rs test_interpolate_1d_input / test_interpolate_2d_input
    for c, expected in _interp_cases():
        assert ora.interpolate(IMG, [c])[0] == expected
    cs = np.array([[0.0, 0.1], [0.1, 0.0]])
    assert_array_equal(ora.interpolate(IMG, cs), [_interp_cases()[2][1], _interp_cases()[3][1]])


def test_interpolation_matches_reference_cpp():
    # oracle/_ref: the reference's own legacy C++ sampler, compiled from its sources by `make ref`
    rng = np.random.default_rng(1)
    img = rng.random((37, 53))
    coords = np.stack([rng.uniform(0, 52, 4000), rng.uniform(0, 36, 4000)], axis=1)
    coords[:500] = np.floor(coords[:500])                 # exact pixels
    coords[500:1000, 0] = np.floor(coords[500:1000, 0])   # integral x only
    coords[1000:1500, 1] = np.floor(coords[1000:1500, 1])  # integral y only
    coords[1500] = [52., 36.]
    ref = ora.ref_bilinear(img, coords)
    if ref is None:
        pytest.skip("oracle/_ref not built (reference sources absent on this machine)")
    assert_array_equal(ora.interpolate(img, coords), ref)


def test_sobel_and_gradient1d():
    # src/gradient.rs test_sobel_x / test_sobel_y / test_gradient1d
    m = np.array([[1., 2., -1., 0.], [0., 0., -1., 1.], [3., -2., 0., -1.], [-2., 1., 1., 2.]])
    gx, gy = ora.sobel(m)
    assert_array_equal(gx, [[0, 0, 0, 0], [0, 7, -1, 0], [0, 4, -4, 0], [0, 0, 0, 0]])
    assert_array_equal(gy, [[0, 0, 0, 0], [0, 5, 3, 0], [0, -2, -6, 0], [0, 0, 0, 0]])
    assert_array_equal(ora.gradient1d([-1., 1., 0., 3., -2.]), [2., -1., 3., -5.])


def test_sobel_is_correlation_not_convolution():
    # src/convolution.rs test_convolve2d: out = sum(kernel * patch), kernel NOT flipped
    rng = np.random.default_rng(2)
    img = rng.random((9, 11))
    gx, gy = ora.sobel(img)
    kx = np.array([[1., 0., -1.], [2., 0., -2.], [1., 0., -1.]])
    ky = np.array([[1., 2., 1.], [0., 0., 0.], [-1., -2., -1.]])
    for y in range(1, 8):
        for x in range(1, 10):
            patch = img[y - 1:y + 2, x - 1:x + 2]
            np.testing.assert_allclose(gx[y, x], (kx * patch).sum(), rtol=0, atol=1e-14)
            np.testing.assert_allclose(gy[y, x], (ky * patch).sum(), rtol=0, atol=1e-14)
    assert not gx[0].any() and not gx[-1].any() and not gx[:, 0].any() and not gx[:, -1].any()


def test_warp():
    # src/warp.rs test_warp_2d / test_warp_1d / test_perspective_warp_1d / _2d
    T = np.array([[0., 0., 1., 0.], [0., 1., 0., 0.], [-1., 0., 0., 4.], [0., 0., 0., 1.]])
    xs1, d1 = ora.warp_vecs(T, [[0., 0.], [2., -1.]], [2., 4.])
    assert_array_equal(xs1, [[0.5, 0.0], [-1.0, 1.0]])
    assert_array_equal(d1, [4., -4.])
    x1, d = ora.warp(T, [0., 0.], 2.)
    assert_array_equal(x1, [0.5, 0.0])
    assert d == 4.
    cam0, cam1 = ora.cam((5., 5.), (20., 30.)), ora.cam((20., 50.), (30., 20.))
    T = np.array([[0., 0., 1., 0.], [0., 1., 0., 0.], [-1., 0., 0., 30.], [0., 0., 0., 1.]])
    u1, d = ora.perspective_warp(T, cam0, cam1, [25., 40.], 10.)
    assert_array_equal(u1, [40., 70.])
    assert d == 20.
    u1, d = ora.perspective_warp(T, cam0, cam1, [0., 10.], 5.)
    assert_array_equal(u1, [32., 0.])
    assert d == 50.


def test_calc_depth0():
    # src/triangulation.rs test_calc_depth0
    R0 = np.array([[0., 0., 1.], [0., 1., 0.], [-1., 0., 0.]])
    R1 = np.array([[0., 0., -1.], [0., 1., 0.], [1., 0., 0.]])
    T_w0, T_w1 = np.eye(4), np.eye(4)
    T_w0[:3, :3], T_w0[:3, 3] = R0, [-3., 0., 1.]
    T_w1[:3, :3], T_w1[:3, 3] = R1, [0., 0., 2.]
    T_0w, _ = ora.mat4_inverse(T_w0)
    T_1w, _ = ora.mat4_inverse(T_w1)
    point = np.array([-1., 0., 1.])
    p0, p1 = ora.transform(T_0w, point), ora.transform(T_1w, point)
    x0, x1 = ora.project(p0), ora.project(p1)
    T_10 = T_1w @ T_w0
    assert ora.calc_depth0(T_10, x0, x1) == p0[2]


# ---------------------------------------------------------------- src/semi_dense/*.rs
def test_calc_ref_ends():
    # src/semi_dense/semi_dense.rs test_calc_ref_ends (also pins the 4x4 inverse)
    T_wk = np.array([[1., 0., 0., -2.], [0., 1., 0., 0.], [0., 0., 1., 7.], [0., 0., 0., 1.]])
    T_wr = np.array([[1., 0., 0., 6.], [0., 1., 0., 0.], [0., 0., 1., 7.], [0., 0., 0., 1.]])
    T_rk = ora.transform_rk(T_wk, T_wr)
    assert_array_equal(T_rk, [[1., 0., 0., -8.], [0., 1., 0., 0.], [0., 0., 1., 0.],
                              [0., 0., 0., 1.]])
    a, b = ora.calc_ref_ends(T_rk, [2., 0.], (2., 3.))
    assert_array_equal(a, [-2., 0.])
    assert_array_equal(b, [-2. / 3., 0.])


def test_step_ratio():
    # src/semi_dense/semi_dense.rs test_step_ratio
    T_rk = np.array([[0., 0., 1., -2.], [0., 1., 0., 2.], [-1., 0., 0., 7.], [0., 0., 0., 1.]])
    flag, ratio = ora.step_ratio(T_rk, [2.0, 2.0], ora.safe_invert(2.0))
    assert flag == 0
    assert abs(ratio - (1. / 2.0) / (1. / 3.0)) < 1e-10
    flag, _ = ora.step_ratio(T_rk, [2.0, 2.0], ora.safe_invert(4.0))
    assert flag == -8


def test_check_us_ref():
    # src/semi_dense/semi_dense.rs test_check_us_ref
    shape = (40, 30)
    assert ora.check_us_ref([[10., 20.], [0., 0.]], 2, shape) == 0
    assert ora.check_us_ref([[10., -2.]], 2, shape) == -5
    assert ora.check_us_ref([[10., -2.], [0., 0.]], 2, shape) == -3
    assert ora.check_us_ref([[0., 0.], [10., -2.]], 2, shape) == -4


def test_calc_key_epipole():
    # src/semi_dense/epipolar.rs test_calc_key_epipole
    T_wk = np.eye(4)
    T_wr = np.array([[1., 0., 0., 3.], [0., 1., 0., 3.], [0., 0., 1., 10.], [0., 0., 0., 1.]])
    assert_array_equal(ora.calc_key_epipole(T_wk, T_wr), [0.3, 0.3])
    T_wk = np.array([[0., 0., 1., 0.], [0., 1., 0., 0.], [-1., 0., 0., 6.], [0., 0., 0., 1.]])
    T_wr = np.array([[0., 0., -1., 6.], [0., 1., 0., 0.], [1., 0., 0., 3.], [0., 0., 0., 1.]])
    assert_array_equal(ora.calc_key_epipole(T_wk, T_wr), [0.5, 0.])
    T_wk = np.array([[-1., 0., 0., 0.], [0., 1., 0., 0.], [0., 0., -1., 4.], [0., 0., 0., 1.]])
    T_wr = np.array([[0., 0., -1., 4.], [0., 1., 0., 0.], [1., 0., 0., 2.], [0., 0., 0., 1.]])
    assert_array_equal(ora.calc_key_epipole(T_wk, T_wr), [-2., 0.])


def test_key_coordinates():
    # src/semi_dense/epipolar.rs test_key_coordinates
    expected = [[7. - 2. * 3., 8. - 2. * 4.], [7. - 1. * 3., 8. - 1. * 4.], [7., 8.],
                [7. + 1. * 3., 8. + 1. * 4.], [7. + 2. * 3., 8. + 2. * 4.]]
    assert_array_equal(ora.key_coordinates([9., 12.], [7., 8.], 5.), expected)


def test_ref_coordinates():
    # src/semi_dense/epipolar.rs test_ref_coordinates
    x_min, x_max = np.array([-15., -20.]), np.array([15., 20.])
    xs = ora.ref_coordinates(x_min, x_max - x_min, 5.0)
    assert_array_equal(xs, [[-15., -20.], [-12., -16.], [-9., -12.], [-6., -8.], [-3., -4.],
                            [0., 0.], [3., 4.], [6., 8.], [9., 12.], [12., 16.]])


def test_intensity_search():
    # src/semi_dense/intensities.rs test_intensity_search (holds under the normalised SSD, F4)
    assert ora.search([-4., 3., 2., 4., -1., 3., 1.], [1., -1., 2.]) == 4
    assert ora.search([-4., 3., 1., -1.], [1., -1.]) == 3
    assert ora.search([1., -1., -4., 3.], [1., -1.]) == 1


def test_intensity_search_first_minimum_wins():
    # intensities.rs:24 `e < min_error` is strict: ties keep the lowest index
    seq = [1., 2., 3., 1., 2., 3., 1., 2., 3.]
    assert ora.search(seq, [1., 2., 3.]) == 0 + 1
    assert ora.search([0., 0., 0., 0., 0., 0.], [1., 2., 3.]) == 0 + 1


def test_depth():
    # src/semi_dense/depth.rs test_calc_inv_depth / test_depth_search_range
    T_rk = np.array([[0., 0., 1., 3.], [0., 1., 0., 2.], [-1., 0., 0., 4.], [0., 0., 0., 1.]])
    assert ora.calc_ref_depth(T_rk, [0.5, 2.0], 4.0) == -2.0 + 4.0
    dmin, dmax = ora.depth_search_range(0.01, 0.10)
    assert dmin == ora.safe_invert(0.10) and dmax == ora.safe_invert(0.01)


def test_geo_var():
    # src/semi_dense/variance.rs test_geo_var
    gradient, direction = np.array([20., -30.]), np.array([6., 2.])
    p = ora.normalize(direction) @ ora.normalize(gradient)
    assert abs(ora.geo_var_(direction, gradient) - 1. / (p * p)) < 1e-12
    assert ora.geo_var_([0., 0.], gradient) == 1. / 1e-16
    assert ora.geo_var_(direction, [0., 0.]) == 1. / 1e-16
    assert ora.geo_var_(direction, [direction[1], -direction[0]]) == 1. / 1e-16


def test_photo_var_and_calc_variance():
    # src/semi_dense/variance.rs test_photo_var / test_calc_variance
    assert ora.photo_var(10.0) == 0.2
    assert ora.calc_variance(0.4, 0.9, 0.8, geo=3., photo=2.) == \
        0.4 * 0.4 * (2. * 2. * 0.8 + 3. * 3. * 0.9)


def test_alpha():
    # src/semi_dense/variance.rs test_alpha_ / test_calc_alpha_
    rot = np.array([[0., -1., 0.], [1., 0., 0.], [0., 0., 1.]])
    t = np.array([2., 4., -3.])
    direction, x_key, x_ref = [0.1, 0.3], np.array([0.3, 0.9]), [-0.6, 0.4]
    y = np.array([0.3, 0.9, 1.])
    for i in (0, 1):
        n = t[i] * (rot[2] @ y) - t[2] * (rot[i] @ y)
        d = t[i] - x_ref[i] * t[2]
        got = ora.alpha_(x_key, x_ref[i], direction[i], rot[i], rot[2], t[i], t[2])
        assert got == direction[i] * n / (d * d)
    T_rk = np.eye(4)
    T_rk[:3, :3], T_rk[:3, 3] = rot, t
    x_ref, _ = ora.warp(T_rk, x_key, 10.0)
    assert ora.calc_alpha_(T_rk, x_key, [0.1, 0.3], 10.0) == \
        ora.alpha_(x_key, x_ref[1], 0.3, rot[1], rot[2], t[1], t[2])
    assert ora.calc_alpha_(T_rk, x_key, [-2., 1.], 10.0) == \
        ora.alpha_(x_key, x_ref[0], -2., rot[0], rot[2], t[0], t[2])


def test_ref_epipolar_direction():
    # src/semi_dense/variance.rs test_ref_epipolar_direction
    T_rk = np.array([[0., 0., 1., -1.], [0., 1., 0., -2.], [-1., 0., 0., 4.], [0., 0., 0., 1.]])
    x_key = [0.4, 1.2]
    d1 = ora.ref_epipolar_direction(T_rk, x_key, (0.5, 1.4))
    a, _ = ora.warp(T_rk, x_key, 0.5)
    b, _ = ora.warp(T_rk, x_key, 1.4)
    d2 = b - a
    assert abs(ora.norm(d1) - 1.) < 1e-15
    assert abs(d1 @ d2 - ora.norm(d2)) < 1e-15


def test_hypothesis():
    # src/semi_dense/hypothesis.rs test_new
    rng_ = (0.4, 1.0)
    for id_, expect in [(0.7, (0.5, 0.9)), (0.3, (0.4, 0.5)), (0.9, (0.7, 1.0))]:
        lo, hi = ora.hypothesis_range(id_, 0.1, rng_)
        assert abs(lo - expect[0]) < 1e-12 and abs(hi - expect[1]) < 1e-12
    for id_ in (0.1, 1.5, 0.2, 1.2):   # incl. the two closed-boundary cases
        assert ora.check_args(id_, 0.1, rng_) == -1
    assert ora.check_args(-0.1, 0.1, rng_) == -7
    assert ora.check_args(0.7, 0.1, rng_) == 0


def test_fusion():
    # src/semi_dense/fusion.rs test_fusion
    mu1 = np.array([[1.9, -2.2], [-3.8, 4.1], [-1.5, 4.5]])
    mu2 = np.array([[-4.1, -2.5], [1.2, 5.0], [6.4, 4.1]])
    v1 = np.array([[4.8, 2.2], [3.1, 6.8], [4.0, 2.1]])
    v2 = np.array([[4.2, 3.1], [0.01, 2.0], [6.0, 3.9]])
    mu0, v0 = ora.fusion_arrays(mu1, mu2, v1, v2)
    # Rust's assert_eq! compares with (v2*m1 + v1*m2); the code computes (m1*v2 + m2*v1): same bits
    assert_array_equal(mu0, (v2 * mu1 + v1 * mu2) / (v1 + v2))
    assert_array_equal(v0, (v1 * v2) / (v1 + v2))


def test_stat():
    # src/semi_dense/stat.rs (its test module never compiles: cfg typo) -> equivalent condition
    rng = np.random.default_rng(3)
    for _ in range(200):
        a, b = rng.uniform(-100, 100, 2)
        v1, v2 = rng.uniform(0, 50, 2)
        c1 = (a - b) * (a - b) <= 4. * v1
        c2 = (a - b) * (a - b) <= 4. * v2
        assert ora.are_statically_same(a, b, v1, v2) == (c1 and c2)


def test_propagate_variance():
    # src/semi_dense/propagation.rs test_propagate_variance
    r = (1. / 2.0) / (1. / 4.0)
    assert ora.propagate_variance(4.0, 2.0, 0.5, 1.0) == (r * r * r * r) * 0.5 + 1.0


def test_propagate():
    # src/semi_dense/propagation.rs test_propagate : 8x8 zoom-out, 16 -> 1 collisions
    W = H = 8
    cam = ora.cam((100., 100.), (W / 2., H / 2.))
    T10 = np.eye(4)
    T10[2, 3] = 300.
    d0, v0 = np.full((H, W), 100.), np.full((H, W), 20.)
    d1, v1 = ora.propagate(T10, cam, cam, d0, v0, 60., 8., 3.)
    exp_d = np.full((H, W), 60.)
    exp_d[3:5, 3:5] = 400.
    np.testing.assert_allclose(d1, exp_d, rtol=0, atol=1e-4)
    var1 = ora.propagate_variance(100., 400., 20., 3.)
    exp_v = np.full((H, W), 8.)
    exp_v[3:5, 3:5] = var1 / 16.
    np.testing.assert_allclose(v1, exp_v, rtol=0, atol=1e-4)


def test_increment_age():
    # src/semi_dense/age.rs test_increment_age
    W, H = 12, 16
    cam = ora.cam((10., 10.), (W / 2., H / 2.))
    T10 = np.eye(4)
    T10[2, 3] = 10.
    age1 = ora.increment_age(np.zeros((H, W), dtype=np.uint64), cam, cam, T10,
                             10.0 * np.ones((H, W)))
    expected = np.zeros((H, W), dtype=np.uint64)
    expected[4:12, 3:9] = 1
    assert_array_equal(age1, expected)


def test_regularize():
    # src/semi_dense/regularization.rs test_regularize (the module is disabled in the reference,
    # F7; this is its only self-consistent vector.  NB: there flag 0 == Success, 1 == "invalid").
    depth = np.array([[1., 2., 4., 2.], [3., 4., 1., 9.], [1., 4., 8., 1.]])
    var = np.array([[1., 4., 3., 5.], [3., 5., 2., 1.], [2., 4., 2., 2.]])
    flag = np.array([[1, 0, 1, 1], [1, 1, 0, 1], [0, 1, 0, 1]], dtype=np.int64)
    out = ora.regularize(depth, var, flag)
    # pixel (0, 2): 3x3 patch rows -1..1, cols 1..3; Success cells are (0,1) and (1,2)
    inv = ora.safe_invert
    num = inv(2.) * inv(4.) + inv(1.) * inv(2.)
    den = inv(4.) + inv(2.)
    assert out[0, 2] == inv(num / den)
    # a pixel with no Success neighbour keeps its depth: make one
    flag2 = np.ones_like(flag)
    assert_array_equal(ora.regularize(depth, var, flag2), depth)
