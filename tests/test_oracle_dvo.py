"""Pins oracle/dvo.py against outputs of the reference's own Python (tests/golden/dvo_golden.npz,
made by tests/golden/make_dvo_golden.py) and checks the restated estimator on a synthetic pair."""
import os

import numpy as np

from oracle import dvo as odvo
from tadataka_b200 import scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dvo_golden.npz"))


def test_jacobian_and_image_gradient_match_reference():
    # tadataka/vo/dvo/jacobian.py
    assert np.array_equal(odvo.calc_jacobian(G["focal"], G["didx"], G["didy"], G["P"]), G["J"])
    DX, DY = odvo.calc_image_gradient(G["image"])
    assert np.array_equal(DX, G["DX"]) and np.array_equal(DY, G["DY"])


def test_robust_weights_match_reference():
    # tadataka/robust/weights.py
    assert np.array_equal(odvo.compute_weights_huber(G["r"]), G["w_huber"])
    assert np.array_equal(odvo.compute_weights_student_t(G["r"]), G["w_student"])
    assert np.array_equal(odvo.compute_weights_tukey(G["r"]), G["w_tukey"])


def test_solver_and_se3_match_reference():
    # tadataka/math.py:34-50, tadataka/se3.py:16-41
    np.testing.assert_allclose(odvo.solve_linear_equation(G["A"], G["b"]), G["x_plain"], rtol=1e-12)
    np.testing.assert_allclose(odvo.solve_linear_equation(G["A"], G["b"], G["w"]), G["x_weighted"],
                               rtol=1e-12)
    for xi, t, Gm in zip(G["xis"], G["ts"], G["Gs"]):
        np.testing.assert_allclose(odvo.exp_se3_t_(xi), t, rtol=1e-13, atol=1e-15)
        np.testing.assert_allclose(odvo.pose_from_se3(xi), Gm, rtol=1e-13, atol=1e-15)


def _pair(seed=9000, H=96, W=128):
    """Two views of a synthetic world with exact depth; small motion."""
    cam = np.array([110.0, 110.0, (W - 1) / 2.0, (H - 1) / 2.0])
    world = scene.World(scene.SET_S, seed)
    T_w0 = scene.pose()
    T_w1 = scene.pose(scene.rotation(0.004, -0.006, 0.003), (1.5, -0.6, 0.9))
    I0, D0 = world.render(T_w0, cam, H, W)
    I1, _ = world.render(T_w1, cam, H, W)
    T10_true = np.linalg.inv(T_w1) @ T_w0
    return cam, I0, D0, I1, T10_true


def test_estimator_reduces_the_photometric_error():
    # the property tests/vo/test_dvo.py:33-44 checks: better than identity, close to the truth
    cam, I0, D0, I1, T10_true = _pair()
    e_id = odvo.photometric_error(np.eye(4), cam, cam, I0, D0, I1)
    e_true = odvo.photometric_error(T10_true, cam, cam, I0, D0, I1)
    for weights in (None, np.ones(I0.shape), "huber", "student-t"):
        T = odvo.estimate_pose_change(cam, cam, I0, D0, I1, np.eye(4), weights, max_iter=20)
        e = odvo.photometric_error(T, cam, cam, I0, D0, I1)
        assert e < e_id, weights
        assert e < max(e_true * 3.0, e_id * 0.5), (weights, e, e_true, e_id)
