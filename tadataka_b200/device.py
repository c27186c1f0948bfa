"""Device-resident, batched form of the semi-dense path: torch CUDA tensors in, torch CUDA
tensors out, no host round trip.  torch is used only for device memory and streams; all
arithmetic happens in libtdk_b200.so (include/tdk.h).

Shapes (P keyframes, R reference frames each):
    key_imgs [P,H,W] f64   key_cams [P,4] f64 {fx,fy,ox,oy}   T_wk [P,4,4] f64
    ref_imgs [P,R,H,W] f64 ref_cams [P,R,4] f64               T_wr [P,R,4,4] f64
    age [P,H,W] u64 (torch.uint64, or int64 with the same bits)   prior maps [P,H,W] f64
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

_F64 = torch.float64
_workspaces = {}


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("tadataka_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _chk(t, name, dtype, shape=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name} must be a CUDA tensor")
    if t.dtype != dtype and not (dtype == torch.uint64 and t.dtype == torch.int64):
        raise TypeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name} has shape {tuple(t.shape)}, expected {tuple(shape)}")
    return t


def workspace(nbytes, device, tag):
    """Grow-only scratch buffer per (device, tag, stream)."""
    key = (device.index, tag, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def params_array(params):
    """6 host doubles {min_depth, max_depth, geo, photo, ref_step, min_gradient}."""
    if hasattr(params, "as_array"):
        params = params.as_array()
    a = np.ascontiguousarray(params, dtype=np.float64)
    if a.shape != (6,):
        raise ValueError("params must be 6 doubles")
    return a


def update_depth(key_imgs, key_cams, T_wk, ref_imgs, ref_cams, T_wr, age, prior_depth,
                 prior_variance, params, debug=False, hist=False, check=True, out=None, fuse=False):
    """Batched update_depth (reference: src/semi_dense/semi_dense.rs:160-234).

    Returns (depth, variance, flag[, argmin][, hist]).  `check=True` reads the status word back
    (one stream synchronisation) and raises ValueError on age/pose/search-length errors.
    `out=(depth, variance, flag)` reuses preallocated output tensors.
    `fuse=True` fuses each successful measurement with its prior inside the kernel (Gaussian
    fusion in inverse depth, fusion.rs:3-11 / propagation.rs:27-34) instead of returning it raw.
    `hist=True` returns a fresh [P,10] int64 count of every flag per keyframe; `hist=<tensor>`
    ADDS the counts to that [P,10] int64 tensor (no allocation).
    """
    _require_cuda()
    _chk(key_imgs, "key_imgs", _F64)
    if key_imgs.dim() != 3:
        raise ValueError("key_imgs must be [P,H,W]")
    P, H, W = key_imgs.shape
    _chk(ref_imgs, "ref_imgs", _F64)
    if ref_imgs.dim() != 4 or ref_imgs.shape[0] != P or tuple(ref_imgs.shape[2:]) != (H, W):
        raise ValueError("ref_imgs must be [P,R,H,W]")
    R = ref_imgs.shape[1]
    _chk(key_cams, "key_cams", _F64, (P, 4))
    _chk(T_wk, "T_wk", _F64, (P, 4, 4))
    _chk(ref_cams, "ref_cams", _F64, (P, R, 4))
    _chk(T_wr, "T_wr", _F64, (P, R, 4, 4))
    _chk(age, "age", torch.uint64, (P, H, W))
    _chk(prior_depth, "prior_depth", _F64, (P, H, W))
    _chk(prior_variance, "prior_variance", _F64, (P, H, W))
    pa = params_array(params)
    dev = key_imgs.device
    with torch.cuda.device(dev):
        if out is None:
            od = torch.empty((P, H, W), dtype=_F64, device=dev)
            ov = torch.empty((P, H, W), dtype=_F64, device=dev)
            of = torch.empty((P, H, W), dtype=torch.int64, device=dev)
        else:
            od, ov, of = out
            _chk(od, "out depth", _F64, (P, H, W))
            _chk(ov, "out variance", _F64, (P, H, W))
            _chk(of, "out flag", torch.int64, (P, H, W))
        dbg = torch.empty((P, H, W), dtype=torch.int32, device=dev) if debug else None
        if isinstance(hist, torch.Tensor):
            hs = _chk(hist, "hist", torch.int64, (P, 10))
        else:
            hs = torch.zeros((P, 10), dtype=torch.int64, device=dev) if hist else None
        nbytes = _lib.update_depth_workspace_bytes(P, R, H, W)
        ws = workspace(nbytes, dev, "update_depth")
        rc = _lib.update_depth_dev(_ptr(key_imgs), _ptr(key_cams), _ptr(T_wk), _ptr(ref_imgs),
                                   _ptr(ref_cams), _ptr(T_wr), _ptr(age), _ptr(prior_depth),
                                   _ptr(prior_variance), pa.ctypes.data_as(C.c_void_p), P, R, H, W,
                                   _ptr(od), _ptr(ov), _ptr(of), _ptr(dbg), _ptr(hs), int(bool(fuse)),
                                   _ptr(ws), ws.numel(), _stream())
        _lib.check(rc)
        if check:
            _lib.check(_lib.update_depth_status(_ptr(ws), _stream()))
    res = [od, ov, of]
    if debug:
        res.append(dbg)
    if hs is not None:
        res.append(hs)
    return tuple(res)


def update_depth_frames(key_img, key_cam, T_wk, ref_imgs, ref_cams, T_wr, age, prior_depth,
                        prior_variance, params, debug=False, fuse=False, device_tables=False):
    """One keyframe against a list of separately allocated reference images (a device-resident
    keyframe ring).  key_img [H,W]; ref_imgs: list of R tensors [H,W]; ref_cams [R,4];
    T_wr [R,4,4]; maps [H,W].  Returns (depth, variance, flag[, argmin]) as [H,W] tensors.
    device_tables: hand the image pointers over as a device array (tdk_update_depth_tables_dev)."""
    _require_cuda()
    _chk(key_img, "key_img", _F64)
    H, W = key_img.shape
    R = len(ref_imgs)
    if R == 0:
        raise ValueError("need at least one reference frame")
    for i, r in enumerate(ref_imgs):
        _chk(r, f"ref_imgs[{i}]", _F64, (H, W))
    _chk(key_cam, "key_cam", _F64, (4,))
    _chk(T_wk, "T_wk", _F64, (4, 4))
    _chk(ref_cams, "ref_cams", _F64, (R, 4))
    _chk(T_wr, "T_wr", _F64, (R, 4, 4))
    _chk(age, "age", torch.uint64, (H, W))
    _chk(prior_depth, "prior_depth", _F64, (H, W))
    _chk(prior_variance, "prior_variance", _F64, (H, W))
    pa = params_array(params)
    dev = key_img.device
    with torch.cuda.device(dev):
        od = torch.empty((H, W), dtype=_F64, device=dev)
        ov = torch.empty((H, W), dtype=_F64, device=dev)
        of = torch.empty((H, W), dtype=torch.int64, device=dev)
        dbg = torch.empty((H, W), dtype=torch.int32, device=dev) if debug else None
        kp = (C.c_void_p * 1)(key_img.data_ptr())
        rp = (C.c_void_p * R)(*[r.data_ptr() for r in ref_imgs])
        ws = workspace(_lib.update_depth_workspace_bytes(1, R, H, W), dev, "update_depth")
        entry = _lib.update_depth_frames_dev
        if device_tables:
            table = torch.tensor([key_img.data_ptr()] + [r.data_ptr() for r in ref_imgs], dtype=torch.int64, device=dev)
            kp, rp = C.c_void_p(table.data_ptr()), C.c_void_p(table.data_ptr() + 8)
            entry = _lib.update_depth_tables_dev
        rc = entry(kp, _ptr(key_cam), _ptr(T_wk), rp, _ptr(ref_cams),
                   _ptr(T_wr), _ptr(age), _ptr(prior_depth),
                   _ptr(prior_variance), pa.ctypes.data_as(C.c_void_p), 1, R,
                   H, W, _ptr(od), _ptr(ov), _ptr(of), _ptr(dbg), None,
                   int(bool(fuse)), _ptr(ws), ws.numel(), _stream())
        _lib.check(rc)
        # the pointer tables are copied with cudaMemcpyAsync from these host arrays: the status
        # read below synchronises the stream before they go out of scope
        _lib.check(_lib.update_depth_status(_ptr(ws), _stream()))
    return (od, ov, of, dbg) if debug else (od, ov, of)


def fusion(mu1, mu2, var1, var2, flag=None, depth_domain=False):
    """Gaussian fusion (reference: src/semi_dense/fusion.rs:13-42).  With `flag`, only elements
    whose flag is SUCCESS are fused; the others keep (mu1, var1).  `depth_domain=True`: mu1/mu2 are
    depths, fused in inverse depth and returned as a depth (propagation.rs:27-34)."""
    _require_cuda()
    shape = mu1.shape
    for t, n in ((mu1, "mu1"), (mu2, "mu2"), (var1, "var1"), (var2, "var2")):
        _chk(t, n, _F64, shape)
    if flag is not None:
        _chk(flag, "flag", torch.int64, shape)
    with torch.cuda.device(mu1.device):
        mu, var = torch.empty_like(mu1), torch.empty_like(mu1)
        _lib.check(_lib.fusion_dev(_ptr(mu1), _ptr(mu2), _ptr(var1), _ptr(var2), _ptr(flag),
                                   mu1.numel(), int(bool(depth_domain)), _ptr(mu), _ptr(var),
                                   _stream()))
    return mu, var


def _as3(t):
    return t.unsqueeze(0) if t.dim() == 2 else t


def regularize(depth, variance, flag):
    """3x3 validity-weighted regularisation (reference: src/semi_dense/regularization.rs:29-64).
    [H,W] or [P,H,W]."""
    _require_cuda()
    d3 = _as3(depth)
    P, H, W = d3.shape
    _chk(depth, "depth", _F64)
    _chk(variance, "variance", _F64, depth.shape)
    _chk(flag, "flag", torch.int64, depth.shape)
    with torch.cuda.device(depth.device):
        out = torch.empty_like(depth)
        _lib.check(_lib.regularize_dev(_ptr(depth), _ptr(variance), _ptr(flag), P, H, W, _ptr(out),
                                       _stream()))
    return out


def increment_age(age0, cam0, cam1, T10, depth0):
    """reference: src/semi_dense/age.rs:6-32.  age0/depth0 [H,W] or [P,H,W]; cams [4] or [P,4];
    T10 [4,4] or [P,4,4]."""
    _require_cuda()
    a3 = _as3(age0)
    P, H, W = a3.shape
    _chk(age0, "age0", torch.uint64)
    _chk(depth0, "depth0", _F64, age0.shape)
    _chk(cam0, "cam0", _F64)
    _chk(cam1, "cam1", _F64)
    _chk(T10, "T10", _F64)
    if cam0.numel() != 4 * P or cam1.numel() != 4 * P or T10.numel() != 16 * P:
        raise ValueError("cam0/cam1/T10 do not match the number of frames")
    dev = age0.device
    with torch.cuda.device(dev):
        out = torch.empty_like(age0)
        ws = workspace(_lib.increment_age_workspace_bytes(P, H, W), dev, "increment_age")
        _lib.check(_lib.increment_age_dev(_ptr(age0), _ptr(cam0), _ptr(cam1), _ptr(T10),
                                          _ptr(depth0), P, H, W, _ptr(out), _ptr(ws), ws.numel(),
                                          _stream()))
    return out


def propagate(T10, cam0, cam1, depth0, variance0, default_depth, default_variance,
              uncertaintity_bias):
    """reference: src/semi_dense/propagation.rs:48-92 (argument order of
    rust_bindings.semi_dense.propagate).  Maps [H,W] or [P,H,W]."""
    _require_cuda()
    d3 = _as3(depth0)
    P, H, W = d3.shape
    _chk(depth0, "depth0", _F64)
    _chk(variance0, "variance0", _F64, depth0.shape)
    _chk(cam0, "cam0", _F64)
    _chk(cam1, "cam1", _F64)
    _chk(T10, "T10", _F64)
    if cam0.numel() != 4 * P or cam1.numel() != 4 * P or T10.numel() != 16 * P:
        raise ValueError("cam0/cam1/T10 do not match the number of frames")
    dev = depth0.device
    with torch.cuda.device(dev):
        d1, v1 = torch.empty_like(depth0), torch.empty_like(depth0)
        ws = workspace(_lib.propagate_workspace_bytes(P, H, W), dev, "propagate")
        _lib.check(_lib.propagate_dev(_ptr(T10), _ptr(cam0), _ptr(cam1), _ptr(depth0),
                                      _ptr(variance0), P, H, W, float(default_depth),
                                      float(default_variance), float(uncertaintity_bias), _ptr(d1),
                                      _ptr(v1), _ptr(ws), ws.numel(), _stream()))
    return d1, v1


def propagate_with_age(T10, cam0, cam1, depth0, variance0, age0, default_depth, default_variance,
                       uncertaintity_bias):
    """increment_age + propagate through one warp of the map (tdk_propagate_age_dev): what the
    reference's loop computes with two calls on the same (T10, cameras, depth0)
    (examples/semi_dense_vo.py:185-194).  -> (depth1, variance1, age1)."""
    _require_cuda()
    d3 = _as3(depth0)
    P, H, W = d3.shape
    _chk(depth0, "depth0", _F64)
    _chk(variance0, "variance0", _F64, depth0.shape)
    _chk(age0, "age0", torch.uint64, depth0.shape)
    _chk(cam0, "cam0", _F64)
    _chk(cam1, "cam1", _F64)
    _chk(T10, "T10", _F64)
    if cam0.numel() != 4 * P or cam1.numel() != 4 * P or T10.numel() != 16 * P:
        raise ValueError("cam0/cam1/T10 do not match the number of frames")
    dev = depth0.device
    with torch.cuda.device(dev):
        d1, v1 = torch.empty_like(depth0), torch.empty_like(depth0)
        a1 = torch.empty_like(age0)
        ws = workspace(_lib.propagate_workspace_bytes(P, H, W), dev, "propagate")
        _lib.check(_lib.propagate_age_dev(_ptr(T10), _ptr(cam0), _ptr(cam1), _ptr(depth0), _ptr(variance0),
                                          _ptr(age0), P, H, W, float(default_depth), float(default_variance),
                                          float(uncertaintity_bias), _ptr(d1), _ptr(v1), _ptr(a1), _ptr(ws),
                                          ws.numel(), _stream()))
    return d1, v1, a1


# ------------------------------------------------------------------ vectorised helpers
def warp_vecs(T10, xs0, depths0):
    _require_cuda()
    n = depths0.numel()
    _chk(T10, "T10", _F64, (4, 4))
    _chk(xs0, "xs0", _F64, (n, 2))
    _chk(depths0, "depths0", _F64, (n,))
    with torch.cuda.device(xs0.device):
        xs1, d1 = torch.empty_like(xs0), torch.empty_like(depths0)
        _lib.check(_lib.warp_vecs_dev(_ptr(T10), _ptr(xs0), _ptr(depths0), n, _ptr(xs1), _ptr(d1),
                                      _stream()))
    return xs1, d1


def interpolate(image, coords):
    """Bilinear samples at (x, y) coordinates; raises ValueError if any lies outside the image
    (as tadataka/interpolation/__init__.py:23-29 does)."""
    _require_cuda()
    _chk(image, "image", _F64)
    H, W = image.shape
    n = coords.shape[0]
    _chk(coords, "coords", _F64, (n, 2))
    with torch.cuda.device(image.device):
        out = torch.empty((n,), dtype=_F64, device=image.device)
        status = torch.zeros((1,), dtype=torch.int32, device=image.device)
        _lib.check(_lib.interpolate_dev(_ptr(image), H, W, _ptr(coords), n, _ptr(out), _ptr(status),
                                        _stream()))
        if int(status.item()) != 0:
            raise ValueError("Coordinates {} out of image range".format("[...]"))
    return out


def project_vecs(ps):
    _require_cuda()
    n = ps.shape[0]
    _chk(ps, "ps", _F64, (n, 3))
    with torch.cuda.device(ps.device):
        xs = torch.empty((n, 2), dtype=_F64, device=ps.device)
        _lib.check(_lib.project_vecs_dev(_ptr(ps), n, _ptr(xs), _stream()))
    return xs


def inv_project_vecs(xs, depths):
    _require_cuda()
    n = xs.shape[0]
    _chk(xs, "xs", _F64, (n, 2))
    _chk(depths, "depths", _F64, (n,))
    with torch.cuda.device(xs.device):
        ps = torch.empty((n, 3), dtype=_F64, device=xs.device)
        _lib.check(_lib.inv_project_vecs_dev(_ptr(xs), _ptr(depths), n, _ptr(ps), _stream()))
    return ps


def transform(T10, ps0):
    _require_cuda()
    n = ps0.shape[0]
    _chk(T10, "T10", _F64, (4, 4))
    _chk(ps0, "ps0", _F64, (n, 3))
    with torch.cuda.device(ps0.device):
        ps1 = torch.empty_like(ps0)
        _lib.check(_lib.transform_dev(_ptr(T10), _ptr(ps0), n, _ptr(ps1), _stream()))
    return ps1


def calc_depth0(T10, xs0, xs1):
    _require_cuda()
    n = xs0.shape[0]
    _chk(T10, "T10", _F64, (4, 4))
    _chk(xs0, "xs0", _F64, (n, 2))
    _chk(xs1, "xs1", _F64, (n, 2))
    with torch.cuda.device(xs0.device):
        out = torch.empty((n,), dtype=_F64, device=xs0.device)
        _lib.check(_lib.calc_depth0_dev(_ptr(T10), _ptr(xs0), _ptr(xs1), n, _ptr(out), _stream()))
    return out


def age_usage(age, max_age):
    """Pixels per age value: age [P,H,W] (or [H,W]) -> int64 tensor [P, max_age + 2] (or
    [max_age + 2]); bin k < max_age + 1 counts age == k, the last bin every older pixel.
    What a caller needs to drop reference frames no pixel points to any more
    (examples/semi_dense_vo.py:202 leaves that as a TODO)."""
    _require_cuda()
    squeeze = age.dim() == 2
    a = age.unsqueeze(0) if squeeze else age
    _chk(a, "age", torch.uint64)
    if a.dim() != 3:
        raise ValueError("age must be [P,H,W] or [H,W]")
    P, H, W = a.shape
    with torch.cuda.device(a.device):
        counts = torch.empty((P, int(max_age) + 2), dtype=torch.int64, device=a.device)
        _lib.check(_lib.age_usage_dev(_ptr(a), P, H, W, int(max_age), _ptr(counts), _stream()))
    return counts[0] if squeeze else counts
