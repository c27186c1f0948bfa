"""GPU parity of the direct pose-change estimator (tadataka_b200.dvo, csrc/dvo.cu) against the
numpy restatement of tadataka/vo/dvo (oracle/dvo.py).  The sums are floating-point reductions in
another order than numpy's, so the comparisons are to a stated tolerance, not bit for bit:
    counts and the image gradient      exact
    J^T W J, J^T W r, error sums       1e-10 relative (to the largest entry)
    pose after the full iteration      1e-7 absolute on the 4x4
"""
import numpy as np
import pytest
import torch

import oracle as ora
from oracle import dvo as odvo
from tadataka_b200 import scene

pytestmark = pytest.mark.gpu
RTOL_SUMS = 1e-10
ATOL_POSE = 1e-7


def _pair(seed=9000, H=96, W=128, t=(1.5, -0.6, 0.9), rot=(0.004, -0.006, 0.003)):
    cam = np.array([110.0, 110.0, (W - 1) / 2.0, (H - 1) / 2.0])
    world = scene.World(scene.SET_S, seed)
    T_w1 = scene.pose(scene.rotation(*rot), t)
    I0, D0 = world.render(scene.pose(), cam, H, W)
    I1, _ = world.render(T_w1, cam, H, W)
    return cam, I0, D0, I1, np.linalg.inv(T_w1)


def _oracle_normal_equations(cam0, cam1, I0, D0, I1, T10, weights):
    """J^T W J, J^T W r, count built the way calc_pose_update builds J, r and the weights."""
    us0 = odvo.image_coordinates(I0.shape)
    P0 = ora.inv_project_vecs(odvo.normalize(cam0, us0.astype(np.float64)), D0.flatten())
    P1 = np.dot(T10[:3, :3], P0.T).T + T10[:3, 3]
    GX1, GY1 = odvo.calc_image_gradient(I1)
    us1 = odvo.unnormalize(cam1, ora.project_vecs(P1))
    mask = odvo.is_in_image_range(us1, I0.shape) & (P1[:, 2] > 0)
    r = (I0 - I1).flatten()[mask]
    J = odvo.calc_jacobian(cam1[0:2], ora.interpolate(GX1, us1[mask]), ora.interpolate(GY1, us1[mask]),
                           P1[mask])
    if weights is None:
        w = np.ones(r.shape)
    elif isinstance(weights, str):
        w = odvo.compute_weights(weights, r)
    else:
        w = weights.flatten()[mask]
    return (J * w[:, None]).T @ J, (J * w[:, None]).T @ r, int(mask.sum())


def test_image_gradient_is_numpy_gradient():
    from tadataka_b200 import dvo
    rng = np.random.default_rng(1)
    for shape in [(13, 17), (4, 64), (97, 131), (2, 2)]:
        img = rng.random(shape)
        gx, gy = dvo.image_gradient(img)
        DX, DY = odvo.calc_image_gradient(img)
        assert np.array_equal(gx.cpu().numpy(), DX) and np.array_equal(gy.cpu().numpy(), DY)


@pytest.mark.parametrize("weights", [None, "array", "huber", "student-t", "tukey"])
def test_normal_equations_and_error_match_oracle(weights):
    from tadataka_b200 import dvo
    cam, I0, D0, I1, T10_true = _pair()
    rng = np.random.default_rng(2)
    wmap = rng.uniform(0.2, 2.0, I0.shape) if weights == "array" else weights
    level = dvo._Level(cam, cam, I0, D0, I1, wmap if isinstance(wmap, np.ndarray) else None)
    for T10 in (np.eye(4), T10_true, scene.pose(scene.rotation(0.01, 0.02, -0.01), (3., 1., -2.))):
        A, b, n = level.normal_equations(T10, wmap)
        oA, ob, on = _oracle_normal_equations(cam, cam, I0, D0, I1, T10, wmap)
        assert n == on
        assert np.max(np.abs(A - oA)) <= RTOL_SUMS * np.max(np.abs(oA))
        assert np.max(np.abs(b - ob)) <= RTOL_SUMS * np.max(np.abs(ob))
        e, oe = level.photometric_error(T10), odvo.photometric_error(T10, cam, cam, I0, D0, I1)
        assert abs(e - oe) <= RTOL_SUMS * oe


def test_tukey_scale_is_the_exact_median_absolute_deviation():
    """tdk_dvo_tukey_scale_dev (radix-select medians on the device) == c * np.median(|r - np.median(r)|)
    of the usable pixels' residuals, bit for bit, for odd and even counts
    (tadataka/robust/weights.py:29-35)."""
    from tadataka_b200 import dvo
    for seed, H, W in ((9100, 60, 80), (9101, 61, 83), (9102, 96, 128)):
        cam, I0, D0, I1, T10_true = _pair(seed=seed, H=H, W=W)
        level = dvo._Level(cam, cam, I0, D0, I1)
        for T10 in (np.eye(4), T10_true):
            us0 = odvo.image_coordinates(I0.shape)
            P0 = ora.inv_project_vecs(odvo.normalize(cam, us0.astype(np.float64)), D0.flatten())
            P1 = np.dot(T10[:3, :3], P0.T).T + T10[:3, 3]
            us1 = odvo.unnormalize(cam, ora.project_vecs(P1))
            mask = odvo.is_in_image_range(us1, I0.shape) & (P1[:, 2] > 0)
            r = (I0 - I1).flatten()[mask]
            want = 1.4826 * np.median(np.abs(r - np.median(r)))
            assert level._tukey_scale(np.ascontiguousarray(T10)) == want, (seed, r.size)


@pytest.mark.parametrize("weights", [None, "huber", "student-t", "tukey"])
def test_single_level_estimator_follows_the_oracle(weights):
    from tadataka_b200 import dvo
    cam, I0, D0, I1, T10_true = _pair()
    gtrace, otrace = [], []
    T = dvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), weights, max_iter=20, trace=gtrace)
    oT = odvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), weights, max_iter=20, trace=otrace)
    assert len(gtrace) == len(otrace) and len(gtrace) >= 2
    assert np.max(np.abs(T - oT)) <= ATOL_POSE
    np.testing.assert_allclose(gtrace[0][0], otrace[0][0], rtol=1e-8, atol=1e-12)   # first update xi
    # and it does what the reference's own test asks (tests/vo/test_dvo.py:33-44)
    e_id = odvo.photometric_error(np.eye(4), cam, cam, I0, D0, I1)
    e_true = odvo.photometric_error(T10_true, cam, cam, I0, D0, I1)
    e = odvo.photometric_error(T, cam, cam, I0, D0, I1)
    assert e < e_id and e < max(3.0 * e_true, 0.5 * e_id)


def test_pyramid_estimator_and_error_cases():
    from rust_bindings.camera import CameraParameters
    from tadataka_b200 import dvo
    cam, I0, D0, I1, T10_true = _pair(seed=9001, H=120, W=160, t=(4.0, -1.5, 2.0), rot=(0.01, -0.015, 0.008))
    cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
    est = dvo.PoseChangeEstimator(cp, cp, n_coarse_to_fine=4, max_iter=20)
    e_id = odvo.photometric_error(np.eye(4), cam, cam, I0, D0, I1)
    e_true = odvo.photometric_error(T10_true, cam, cam, I0, D0, I1)
    for weights in (None, np.ones(I0.shape), "huber", "student-t", "tukey"):
        T = est(I0, D0, I1, weights)
        e = odvo.photometric_error(T, cam, cam, I0, D0, I1)
        assert e < e_id, weights
        # the biweight cuts off every residual beyond 4.7 sigma_MAD, so on this pair it stops earlier
        bound = 0.5 * e_id if isinstance(weights, str) and weights == "tukey" else max(2.0 * e_true, 0.25 * e_id)
        assert e < bound, (weights, e, e_true, e_id)
    with pytest.raises(ValueError, match="No such weights 'random'"):
        est(I0, D0, I1, "random")
    # a pose that throws every pixel out of the image: the reference warns and returns its input
    far = scene.pose(None, (1e6, 0., 0.))
    with pytest.warns(RuntimeWarning, match="too large"):
        out = dvo.estimate_pose_change(cam, cam, I0, D0, I1, far, None, max_iter=3)
    assert np.array_equal(out, far)


@pytest.mark.parametrize("weights", [None, "huber", "map"])
def test_device_resident_loop_equals_the_host_loop_and_the_oracle(weights):
    """tdk_dvo_iterate_dev (solve, exponential and accept / stop rule on the device) against the
    per-iteration host loop of the same kernels (trace given) and against the numpy oracle."""
    from tadataka_b200 import dvo
    # a motion large enough for ten or so accepted steps: every one of them reads the pose the
    # previous one left in device memory (a pose element dropped there shows up as 1e-2 in the result)
    cam, I0, D0, I1, _ = _pair(seed=9100, H=240, W=320, t=(6.0, -2.5, 3.5), rot=(0.012, -0.018, 0.009))
    if weights == "map":
        rng = np.random.default_rng(4)
        weights = rng.uniform(0.2, 1.0, size=I0.shape)
    trace = []
    T_host = dvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), weights, max_iter=20, trace=trace)
    T_dev = dvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), weights, max_iter=20)
    oT = odvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), weights, max_iter=20)
    assert np.max(np.abs(T_dev - T_host)) <= 1e-9
    assert np.max(np.abs(T_dev - oT)) <= ATOL_POSE
    # the loop's own bookkeeping: as many accepted steps as the host loop took before its error rose
    level = dvo._Level(cam, cam, I0, D0, I1, weights if isinstance(weights, np.ndarray) else None)
    import torch
    from tadataka_b200 import _lib, device
    state = torch.empty(128, dtype=torch.float64, device="cuda")
    T0 = np.eye(4)
    mode = 2 if isinstance(weights, str) else (1 if weights is not None else 0)
    _lib.check(_lib.dvo_iterate_dev(device._ptr(level.I0), device._ptr(level.D0), device._ptr(level.I1),
                                    device._ptr(level.GX1), device._ptr(level.GY1), device._ptr(level.W0), mode,
                                    dvo.HUBER_K if mode == 2 else 0.0, dvo._vp(level.cam0), dvo._vp(level.cam1),
                                    dvo._vp(T0), level.H, level.W, 20, 0, device._ptr(state), device._stream()))
    st = state.cpu().numpy()
    accepted_host = len(trace) - 1 if len(trace) < 20 or trace[-1][1] > trace[-2][1] else len(trace)
    assert int(st[66]) == 0 and int(st[68]) == accepted_host and accepted_host >= 4
    # queued all at once (here) or in batches with resume (estimate_pose_change): the same walk
    assert np.max(np.abs(st[:16].reshape(4, 4) - T_dev)) <= 1e-9 and st[67] == 1.0
    # zero iterations: the start pose comes back untouched
    _lib.check(_lib.dvo_iterate_dev(device._ptr(level.I0), device._ptr(level.D0), device._ptr(level.I1),
                                    device._ptr(level.GX1), device._ptr(level.GY1), device._ptr(level.W0), mode,
                                    dvo.HUBER_K if mode == 2 else 0.0, dvo._vp(level.cam0), dvo._vp(level.cam1),
                                    dvo._vp(T_dev), level.H, level.W, 0, 0, device._ptr(state), device._stream()))
    assert np.array_equal(state.cpu().numpy()[:16].reshape(4, 4), T_dev)


def test_device_pyramid_is_the_numpy_pyramid():
    """rescale_device (Gaussian pre-filter + bilinear resampling on the device) against `rescale`
    (scipy gaussian_filter + map_coordinates): equal to rounding, on ragged and tiny images too; and
    the coarse-to-fine estimator built on it against the same estimator fed a numpy-made pyramid."""
    from tadataka_b200 import dvo
    rng = np.random.default_rng(12)
    for shape in [(97, 131), (480, 640), (7, 5), (33, 2)]:
        img = rng.random(shape)
        for level in range(5):
            scale = 1 / pow(1.5, level)
            if min(int(round(shape[0] * scale)), int(round(shape[1] * scale))) < 1:
                continue
            got = dvo.rescale_device(torch.from_numpy(img).cuda(), scale).cpu().numpy()
            want = dvo.rescale(img, scale)
            assert got.shape == want.shape
            assert np.max(np.abs(got - want)) <= 1e-13, (shape, level)
    cam, I0, D0, I1, _ = _pair(seed=9001, H=120, W=160, t=(4.0, -1.5, 2.0), rot=(0.01, -0.015, 0.008))
    from rust_bindings.camera import CameraParameters
    cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
    T = dvo.PoseChangeEstimator(cp, cp, n_coarse_to_fine=4, max_iter=20)(I0, D0, I1)
    T_np = np.eye(4)
    for level in reversed(range(4)):           # the same walk with the pyramid made by numpy
        scale = 1 / pow(1.5, level)
        lv = dvo._Level(cam * scale, cam * scale, dvo.rescale(I0, scale), dvo.rescale(D0, scale), dvo.rescale(I1, scale))
        T_np = dvo._iterate(lv, T_np, None, 20)
    assert np.max(np.abs(T - T_np)) <= 1e-8
