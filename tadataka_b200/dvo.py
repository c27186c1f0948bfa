"""Direct (photometric) pose-change estimation on the GPU: the reference's
`tadataka.vo.dvo.PoseChangeEstimator` (tadataka/vo/dvo/__init__.py:76-150), the caller on the other
side of the semi-dense keyframe loop (examples/semi_dense_vo.py:45-54 feeds it the depth and
variance maps `update_depth` produces).

Per Gauss-Newton iteration the two per-pixel passes -- building J^T W J / J^T W r
(calc_pose_update, :46-73) and re-evaluating the photometric error (tadataka/metric.py:13-39) --
run as CUDA reductions (csrc/dvo.cu behind include/tdk.h).  For weights None / map / "huber" the
whole loop -- 6x6 solve, SE(3) exponential (tadataka/se3.py:16-31), accept/stop rule (:98-108) --
stays on the device (tdk_dvo_iterate_dev, one read-back per resolution); for "student-t" / "tukey",
or when a trace is asked for, 30 + 2 doubles come back per iteration and those steps run on the host.  Poses are 4x4 matrices T10 (frame 1 <- frame 0) instead of the
reference's `Pose` objects.

Deviations from the reference:
  * the least-squares step solves the 6x6 normal equations (minimum-norm, via lstsq) instead of
    running lstsq on the [N,6] system (tadataka/math.py:34-50): same minimiser, different rounding;
  * the image pyramid (:139-148) uses `rescale` below (Gaussian pre-filter + bilinear; made on the
    device by `rescale_device`), not skimage.transform.rescale, which is not available here;
"""
import ctypes as C
import warnings

import numpy as np
import torch
from scipy.ndimage import gaussian_filter, map_coordinates
from scipy.spatial.transform import Rotation

from . import _lib, device
from .camera import camera_vector

_F64 = torch.float64
HUBER_K = 1.345       # tadataka/robust/weights.py:38
STUDENT_NU = 5.0      # tadataka/robust/weights.py:4
TUKEY_C = 1.4826      # tadataka/robust/weights.py:32 (beta = 4.6851 lives in csrc/dvo.cu)
STUDENT_ITERS = 10


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def _dev(a):
    if isinstance(a, torch.Tensor):
        return device._chk(a, "image", _F64)
    a = np.ascontiguousarray(a)
    if a.dtype != np.float64 or a.ndim != 2:
        raise TypeError("images and maps must be 2-D float64")
    device._require_cuda()
    return torch.from_numpy(a).cuda()


def _skew(w):
    wx, wy, wz = float(w[0]), float(w[1]), float(w[2])
    return np.array([[0., -wz, wy], [wz, 0., -wx], [-wy, wx, 0.]])


def exp_se3(xi):
    """4x4 matrix of the twist xi = (v, w): rotation exp([w]x), translation V(w) v, the map
    tadataka/pose.py:44-46 + tadataka/se3.py:16-31 compute.  Written on the un-normalised skew
    matrix W = [w]x with the closed-form coefficients
        a = sin(t)/t,  b = (1 - cos t)/t^2,  c = (t - sin t)/t^3,   t = |w|:
        R = I + a W + b W^2,   V = I + b W + c W^2,
    and their Taylor series below t = 1e-3 (where the closed forms cancel).  The reference switches to
    its series only below t = 1e-16, so around t ~ 1e-8 its (1 - cos t)/t has cancelled to zero and it
    is off by up to t/2 |v|; this form is not (agreement elsewhere: 1e-13)."""
    xi = np.asarray(xi, dtype=np.float64)
    W = _skew(xi[3:])
    W2 = W @ W
    t2 = float(xi[3] * xi[3] + xi[4] * xi[4] + xi[5] * xi[5])
    if t2 < 1e-6:
        a = 1. - t2 / 6. * (1. - t2 / 20.)
        b = 0.5 - t2 / 24. * (1. - t2 / 30.)
        c = 1. / 6. - t2 / 120. * (1. - t2 / 42.)
    else:
        t = np.sqrt(t2)
        st, ct = np.sin(t), np.cos(t)
        a, b, c = st / t, (1. - ct) / t2, (t - st) / (t2 * t)
    T = np.eye(4)
    T[:3, :3] += a * W + b * W2
    T[:3, 3] = xi[:3] + b * (W @ xi[:3]) + c * (W2 @ xi[:3])
    return T


def image_gradient(image):
    """(DX, DY) = np.gradient along x and y (tadataka/vo/dvo/jacobian.py:27-29), device tensors."""
    image = _dev(image)
    H, W = image.shape
    gx, gy = torch.empty_like(image), torch.empty_like(image)
    _lib.check(_lib.image_gradient_dev(device._ptr(image), H, W, device._ptr(gx), device._ptr(gy),
                                       device._stream()))
    return gx, gy


class _Level:
    """Device-resident inputs of one pyramid level."""

    def __init__(self, cam0, cam1, I0, D0, I1, W0=None):
        self.cam0 = np.ascontiguousarray(cam0, dtype=np.float64)
        self.cam1 = np.ascontiguousarray(cam1, dtype=np.float64)
        self.I0, self.D0, self.I1 = _dev(I0), _dev(D0), _dev(I1)
        if not (self.I0.shape == self.D0.shape == self.I1.shape):
            raise ValueError("I0, D0 and I1 must have the same shape")
        self.W0 = _dev(W0) if W0 is not None else None
        if self.W0 is not None and self.W0.shape != self.I0.shape:
            raise ValueError("the weight map must have the image's shape")
        self.H, self.W = self.I0.shape
        self.GX1, self.GY1 = image_gradient(self.I1)
        self.acc = torch.empty(32, dtype=_F64, device=self.I0.device)

    def photometric_error(self, T10):
        T10 = np.ascontiguousarray(T10, dtype=np.float64)
        _lib.check(_lib.dvo_photometric_error_dev(
            device._ptr(self.I0), device._ptr(self.D0), device._ptr(self.I1), _vp(self.cam0),
            _vp(self.cam1), _vp(T10), self.H, self.W, device._ptr(self.acc), device._stream()))
        s, n = self.acc[:2].cpu().numpy()
        return s / n if n > 0 else np.nan      # np.mean of an empty array is nan, as in metric.py

    def _student_t_variance(self, T10):
        variance = 1.0
        for _ in range(STUDENT_ITERS):
            _lib.check(_lib.dvo_student_t_dev(
                device._ptr(self.I0), device._ptr(self.D0), device._ptr(self.I1), variance, STUDENT_NU,
                _vp(self.cam0), _vp(self.cam1), _vp(T10), self.H, self.W, device._ptr(self.acc),
                device._stream()))
            s, n = self.acc[:2].cpu().numpy()
            if n == 0:
                return None
            variance = s / n
        return variance

    def _tukey_scale(self, T10):
        """c * MAD of the residuals of the usable pixels (robust/weights.py:29-35), found on the device."""
        nbytes = _lib.dvo_tukey_workspace_bytes(self.H, self.W)
        ws = device.workspace(nbytes, self.I0.device, "dvo_tukey")
        _lib.check(_lib.dvo_tukey_scale_dev(
            device._ptr(self.I0), device._ptr(self.D0), device._ptr(self.I1), TUKEY_C, _vp(self.cam0),
            _vp(self.cam1), _vp(T10), self.H, self.W, device._ptr(self.acc), device._ptr(ws), ws.numel(),
            device._stream()))
        scale = float(self.acc[0].item())
        return None if np.isnan(scale) else scale

    def normal_equations(self, T10, weights):
        """-> (A [6,6], b [6], n_used) of calc_pose_update, or None if no pixel is usable."""
        T10 = np.ascontiguousarray(T10, dtype=np.float64)
        mode, param, wmap = 0, 0.0, None
        if weights is None:
            pass
        elif isinstance(weights, str):
            if weights == "huber":
                mode, param = 2, HUBER_K
            elif weights == "student-t":
                mode, param = 3, self._student_t_variance(T10)
                if param is None:
                    return None
            elif weights == "tukey":
                mode, param = 4, self._tukey_scale(T10)
                if param is None:
                    return None
            else:
                raise ValueError(f"No such weights '{weights}'")
        else:
            mode, wmap = 1, self.W0
        _lib.check(_lib.dvo_normal_equations_dev(
            device._ptr(self.I0), device._ptr(self.D0), device._ptr(self.I1), device._ptr(self.GX1),
            device._ptr(self.GY1), device._ptr(wmap), mode, float(param), STUDENT_NU, _vp(self.cam0),
            _vp(self.cam1), _vp(T10), self.H, self.W, device._ptr(self.acc), device._stream()))
        acc = self.acc[:30].cpu().numpy()
        if acc[27] == 0:
            return None
        A = np.zeros((6, 6))
        A[np.triu_indices(6)] = acc[:21]
        A = A + np.triu(A, 1).T
        return A, acc[21:27].copy(), int(acc[27])


def estimate_pose_change(camera_params0, camera_params1, I0, D0, I1, T10=None, weights=None,
                         max_iter=20, trace=None):
    """_PoseChangeEstimator.__call__ (tadataka/vo/dvo/__init__.py:76-110) at one resolution.
    I0, D0, I1 (and a weight map) are float64 [H,W] numpy arrays or CUDA tensors; returns T10."""
    cam0 = camera_params0 if isinstance(camera_params0, np.ndarray) else camera_vector(camera_params0)
    cam1 = camera_params1 if isinstance(camera_params1, np.ndarray) else camera_vector(camera_params1)
    wmap = None if weights is None or isinstance(weights, str) else weights
    level = _Level(cam0, cam1, I0, D0, I1, wmap)
    return _iterate(level, np.eye(4) if T10 is None else np.array(T10, dtype=np.float64), weights,
                    max_iter, trace)


def _iterate_on_device(level, T10, weights, max_iter):
    """The whole Gauss-Newton loop in one C-ABI call (tdk_dvo_iterate_dev): normal equations, 6x6
    Cholesky solve, SE(3) exponential, error of the candidate and the accept / stop rule all run on
    the device; one read-back at the end.  For weights None / a weight map / "huber" (student-t and
    tukey need a scale per iteration that is found with host logic).  Returns T10, or None when the
    normal equations were not positive definite (the host loop then decides with its lstsq)."""
    mode, param, wmap = 0, 0.0, None
    if isinstance(weights, str):
        mode, param = 2, HUBER_K
    elif weights is not None:
        mode, wmap = 1, level.W0
    T10 = np.ascontiguousarray(T10, dtype=np.float64)
    state = torch.empty(128, dtype=_F64, device=level.I0.device)
    # iterations are queued in growing batches (4, 8, 16, ...): a queued iteration costs four launches even
    # after the loop has stopped, a look at the state costs one small read-back
    queued, batch, st = 0, 4, None
    while True:
        n = min(batch, max_iter - queued)
        _lib.check(_lib.dvo_iterate_dev(
            device._ptr(level.I0), device._ptr(level.D0), device._ptr(level.I1), device._ptr(level.GX1),
            device._ptr(level.GY1), device._ptr(wmap), mode, float(param), _vp(level.cam0), _vp(level.cam1), _vp(T10),
            level.H, level.W, int(n), int(queued > 0), device._ptr(state), device._stream()))
        queued += n
        st = state.cpu().numpy()
        if st[67] != 0. or queued >= max_iter:
            break
        batch *= 2
    status = int(st[66])
    if status == 2:
        return None
    if status == 1:
        warnings.warn("Camera pose change is too large.", RuntimeWarning)
    return st[:16].reshape(4, 4).copy()


def _iterate(level, T10, weights, max_iter, trace=None):
    if trace is None and (weights is None or not isinstance(weights, str) or weights == "huber"):
        T = _iterate_on_device(level, T10, weights, max_iter)
        if T is not None:
            return T
    prev_error = level.photometric_error(T10)
    for _ in range(max_iter):
        ne = level.normal_equations(T10, weights)
        if ne is None:
            warnings.warn("Camera pose change is too large.", RuntimeWarning)
            return T10
        A, b, _ = ne
        xi = np.linalg.lstsq(A, b, rcond=None)[0]
        candidate = exp_se3(xi) @ T10                  # pose.py:51-53: dpose * pose10
        curr_error = level.photometric_error(candidate)
        if trace is not None:
            trace.append((xi, curr_error))
        if curr_error > prev_error:
            break
        prev_error = curr_error
        T10 = candidate
    return T10


def rescale(image, scale):
    """Anti-aliased bilinear resampling by `scale` (<= 1).  Stands in for skimage.transform.rescale
    (tadataka/vo/dvo/__init__.py:143-148); NOT numerically identical to it."""
    image = np.asarray(image, dtype=np.float64)
    if scale == 1:
        return image.copy()
    H, W = image.shape
    h, w = int(round(H * scale)), int(round(W * scale))
    sigma = max(0.0, (1.0 / scale - 1.0) / 2.0)
    smooth = gaussian_filter(image, sigma, mode="reflect") if sigma > 0 else image
    ys = (np.arange(h) + 0.5) / scale - 0.5
    xs = (np.arange(w) + 0.5) / scale - 0.5
    yy, xx = np.meshgrid(ys, xs, indexing="ij")
    return map_coordinates(smooth, [yy, xx], order=1, mode="reflect")


def rescale_device(image, scale):
    """`rescale` on the device (tdk_gaussian_blur_dev + tdk_resample_bilinear_dev): the same Gaussian
    weights (scipy's: exp(-x^2 / 2 sigma^2) over |x| <= int(4 sigma + 0.5), normalised), the same
    half-sample reflection and sampling grid; agrees with `rescale` to rounding (the sums run in
    another order).  image: CUDA tensor [H,W] f64 -> CUDA tensor [round(H scale), round(W scale)]."""
    image = _dev(image)
    if scale == 1:
        return image.clone()
    H, W = image.shape
    h, w = int(round(H * scale)), int(round(W * scale))
    sigma = max(0.0, (1.0 / scale - 1.0) / 2.0)
    smooth = image
    if sigma > 0:
        radius = int(4.0 * sigma + 0.5)
        if radius > 32:
            raise ValueError("scale too small for the device pyramid (Gaussian radius > 32)")
        x = np.arange(-radius, radius + 1)
        k = np.exp(-0.5 / (sigma * sigma) * x ** 2)
        k = np.ascontiguousarray(k / k.sum())
        tmp, smooth = torch.empty_like(image), torch.empty_like(image)
        _lib.check(_lib.gaussian_blur_dev(device._ptr(image), H, W, _vp(k), radius, device._ptr(tmp),
                                          device._ptr(smooth), device._stream()))
    out = torch.empty((h, w), dtype=_F64, device=image.device)
    _lib.check(_lib.resample_bilinear_dev(device._ptr(smooth), H, W, float(scale), device._ptr(out), h, w,
                                          device._stream()))
    return out


class PoseChangeEstimator:
    """`PoseChangeEstimator(camera_params0, camera_params1, n_coarse_to_fine=5, max_iter=20,
    layer_size_ratio=1.5)`; call it with `(I0, D0, I1, weights=None, pose10=identity)`
    (tadataka/vo/dvo/__init__.py:113-150).  camera_params*: objects with .focal_length / .offset."""

    def __init__(self, camera_params0, camera_params1, n_coarse_to_fine=5, max_iter=20,
                 layer_size_ratio=1.5):
        self.cam0 = camera_vector(camera_params0)
        self.cam1 = camera_vector(camera_params1)
        self.n_coarse_to_fine = n_coarse_to_fine
        self.max_iter = max_iter
        self.layer_size_ratio = layer_size_ratio

    def __call__(self, I0, D0, I1, weights=None, pose10=None):
        """I0, D0, I1 (and a weight map): float64 [H,W] numpy arrays or CUDA tensors.  The images go to
        the device once; every pyramid level is made there (`rescale_device`)."""
        def up(a):
            return _dev(a if isinstance(a, torch.Tensor) else np.asarray(a, dtype=np.float64))
        if not (tuple(I0.shape) == tuple(D0.shape) == tuple(I1.shape)) or len(I0.shape) != 2:
            raise ValueError("I0, D0, I1 must be 2-D arrays of one shape")
        I0, D0, I1 = up(I0), up(D0), up(I1)
        wmap = None
        if weights is not None and not isinstance(weights, str):
            wmap = up(weights)
        T10 = np.eye(4) if pose10 is None else np.array(pose10, dtype=np.float64)
        for level in reversed(range(self.n_coarse_to_fine)):
            scale = 1 / pow(self.layer_size_ratio, level)      # level_to_scale, :42-43
            W0 = rescale_device(wmap, scale) if wmap is not None else None
            lv = _Level(self.cam0 * scale, self.cam1 * scale, rescale_device(I0, scale), rescale_device(D0, scale),
                        rescale_device(I1, scale), W0)         # camera.resize, camera/model.py:69-74
            T10 = _iterate(lv, T10, W0 if W0 is not None else weights, self.max_iter)
        return T10
