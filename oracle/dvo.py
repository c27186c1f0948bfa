"""CPU restatement (numpy) of the reference's direct pose-change estimator, SURVEY 8f rank 1.

TEST INFRASTRUCTURE ONLY, like the rest of oracle/: imported by tests/ (and tools/) as the checker
for tadataka_b200.dvo, never by the product.

Follows, function by function (paths relative to /root/reference):
    tadataka/vo/dvo/__init__.py:46-110   calc_pose_update, _PoseChangeEstimator.__call__
    tadataka/vo/dvo/jacobian.py:8-29     calc_jacobian, calc_image_gradient
    tadataka/metric.py:8-39              photometric_error, PhotometricError
    tadataka/robust/weights.py:4-43      huber / student-t / tukey weights
    tadataka/math.py:34-50               solve_linear_equation (lstsq of the sqrt-weighted system)
    tadataka/se3.py:16-31, so3.py:27-41  exp_se3_t_, tangent_so3 ; tadataka/pose.py:44-53
    tadataka/coordinates.py:7-20, utils.py:35-42, warp.py:78-88, rigid_transform.py:115-123
The geometry primitives (warp, pi, bilinear interpolation) are the ones of the semi-dense oracle
(oracle/tdk_oracle.c), because the reference's Python calls the same Rust functions for them.

Parity status: calc_jacobian, the weight functions and exp_se3_t_ are pinned by golden vectors
produced by executing the reference's own files (tests/golden/make_dvo_golden.py ->
tests/golden/dvo_golden.npz).  The iteration as a whole is NOT pinned against the reference: its
module imports skimage (absent here) at top level, and its image pyramid is skimage's `rescale`,
which is not restated -- the estimator is checked at a single level only ("pyramid unpinned").
"""
import numpy as np
from scipy.spatial.transform import Rotation

import oracle as ora


# ---------------------------------------------------------------------------------- small helpers
def image_coordinates(shape):
    """tadataka/coordinates.py:7-20 : [[x0,y0],[x1,y0],...] in raster order."""
    h, w = shape[0:2]
    xs, ys = np.meshgrid(np.arange(w), np.arange(h))
    return np.column_stack((xs.flatten(), ys.flatten()))


def is_in_image_range(us, shape):
    """tadataka/utils.py:35-42"""
    h, w = shape[0:2]
    return (0 <= us[:, 0]) & (us[:, 0] <= w - 1) & (0 <= us[:, 1]) & (us[:, 1] <= h - 1)


def normalize(cam, us):
    return (us - cam[2:4]) / cam[0:2]


def unnormalize(cam, xs):
    return xs * cam[0:2] + cam[2:4]


def calc_image_gradient(image):
    """tadataka/vo/dvo/jacobian.py:27-29"""
    DY, DX = np.gradient(image)
    return DX, DY


def calc_jacobian(focal_length, didx, didy, P):
    """d(intensity)/d(twist) per point, [N,6]: Kerl's formula as tadataka/vo/dvo/jacobian.py:8-24
    evaluates it (same operations in the same order, so the golden vectors match bit for bit)."""
    a = focal_length[0] * didx          # fx * dI/dx
    b = focal_length[1] * didy          # fy * dI/dy
    X, Y, Z = P[:, 0], P[:, 1], P[:, 2]
    ZZ = Z * Z
    XY = X * Y
    cols = [a / Z,
            b / Z,
            -(a * X + b * Y) / (Z * Z),
            -(a * XY + b * (ZZ + Y * Y)) / ZZ,
            (a * (ZZ + X * X) + b * XY) / ZZ,
            (-a * Y + b * X) / Z]
    return np.stack(cols, axis=1)


# ---------------------------------------------------------------------------------- robust weights
def compute_weights_student_t(r, nu=5, n_iter=10):
    """tadataka/robust/weights.py:4-20: fixed point of the t-distribution scale, then the SQUARE
    ROOT of the t-weights is what comes back."""
    sq = np.power(r, 2)
    var = 1.0
    for _ in range(n_iter):
        var = np.mean(sq * ((nu + 1) / (nu + sq / var)))
    return np.sqrt((nu + 1) / (nu + sq / var))


def compute_weights_huber(r, k=1.345):
    """tadataka/robust/weights.py:38-43: 1 inside the band, k/|r| outside."""
    mag = np.abs(r)
    out = np.ones(r.shape)
    outside = mag > k
    out[outside] = k / mag[outside]
    return out


def compute_weights_tukey(r, beta=4.6851, c=1.4826):
    """tadataka/robust/weights.py:23-35: biweight of r / (c * MAD)."""
    mad = np.median(np.abs(r - np.median(r)))
    u = r / (c * mad)
    out = np.zeros(u.shape)
    inside = np.abs(u) <= beta
    out[inside] = np.power(1 - np.power(u[inside] / beta, 2), 2)
    return out


_WEIGHTS = {"tukey": compute_weights_tukey, "student-t": compute_weights_student_t,
            "huber": compute_weights_huber}


def compute_weights(name, residuals):
    """tadataka/vo/dvo/__init__.py:32-39"""
    if name not in _WEIGHTS:
        raise ValueError(f"No such weights '{name}'")
    return _WEIGHTS[name](residuals)


def solve_linear_equation(A, b, weights=None):
    """tadataka/math.py:34-50 with the default "lstsq" solver: rows scaled by sqrt(weights)."""
    if weights is not None:
        root = np.sqrt(weights)
        A, b = A * root[:, None], b * root
    return np.linalg.lstsq(A, b, rcond=None)[0]


# ---------------------------------------------------------------------------------- SE(3)
def tangent_so3(v):
    """tadataka/so3.py:27-38"""
    return np.array([[0., -v[2], v[1]], [v[2], 0., -v[0]], [-v[1], v[0], 0.]])


def exp_se3_t_(xi):
    """tadataka/se3.py:16-31"""
    v, rotvec = xi[:3], xi[3:]
    theta = np.linalg.norm(rotvec)
    omega = np.zeros(3) if theta == 0 else rotvec / theta
    K = tangent_so3(omega)
    I = np.eye(3)
    if theta < 1e-16:
        V = I + K * theta / 2 + np.dot(K, K) * pow(theta, 2) / 6
    else:
        V = (I + (1 - np.cos(theta)) / theta * K + (theta - np.sin(theta)) / theta * np.dot(K, K))
    return np.dot(V, v)


def pose_from_se3(xi):
    """tadataka/pose.py:44-46 as a 4x4 matrix."""
    T = np.eye(4)
    T[:3, :3] = Rotation.from_rotvec(xi[3:]).as_matrix()
    T[:3, 3] = exp_se3_t_(xi)
    return T


# ---------------------------------------------------------------------------------- the estimator
def photometric_error(T10, cam0, cam1, I0, D0, I1):
    """tadataka/metric.py:13-39 : mean squared difference between I0 and I1 warped by (T10, D0),
    over the pixels whose warped position falls inside the image (no sign test on the depth)."""
    us0 = image_coordinates(D0.shape)
    xs1, _ = ora.warp_vecs(T10, normalize(cam0, us0.astype(np.float64)), D0.flatten())
    us1 = unnormalize(cam1, xs1)
    mask = is_in_image_range(us1, D0.shape)
    i0 = I0[us0[mask, 1], us0[mask, 0]]
    i1 = ora.interpolate(I1, us1[mask])
    d = i0 - i1
    return np.mean(d * d)


def calc_pose_update(cam1, residuals, GX1, GY1, P1, weights):
    """tadataka/vo/dvo/__init__.py:46-73 ; returns xi (6,) or None."""
    us1 = unnormalize(cam1, ora.project_vecs(P1))
    mask = is_in_image_range(us1, GX1.shape) & (P1[:, 2] > 0)
    if not np.any(mask):
        return None
    r = residuals[mask]
    p1 = P1[mask]
    gx1 = ora.interpolate(GX1, us1[mask])
    gy1 = ora.interpolate(GY1, us1[mask])
    J = calc_jacobian(cam1[0:2], gx1, gy1, p1)
    if weights is None:
        return solve_linear_equation(J, r)
    if isinstance(weights, str):
        return solve_linear_equation(J, r, compute_weights(weights, r))
    return solve_linear_equation(J, r, weights.flatten()[mask])


def estimate_pose_change(cam0, cam1, I0, D0, I1, T10, weights=None, max_iter=20, trace=None):
    """tadataka/vo/dvo/__init__.py:76-110 (_PoseChangeEstimator.__call__) on 4x4 matrices.
    `trace`, if a list, receives (xi, candidate_error) per iteration."""
    us0 = image_coordinates(I0.shape)
    xs0 = normalize(cam0, us0.astype(np.float64))
    P0 = ora.inv_project_vecs(xs0, D0.flatten())
    GX1, GY1 = calc_image_gradient(I1)
    residuals = (I0 - I1).flatten()
    prev_error = photometric_error(T10, cam0, cam1, I0, D0, I1)
    for _ in range(max_iter):
        P1 = np.dot(T10[:3, :3], P0.T).T + T10[:3, 3]      # rigid_transform.py:115-123
        xi = calc_pose_update(cam1, residuals, GX1, GY1, P1, weights)
        if xi is None:
            return T10
        candidate = pose_from_se3(xi) @ T10                  # pose.py:51-53
        curr_error = photometric_error(candidate, cam0, cam1, I0, D0, I1)
        if trace is not None:
            trace.append((xi, curr_error))
        if curr_error > prev_error:
            break
        prev_error = curr_error
        T10 = candidate
    return T10
