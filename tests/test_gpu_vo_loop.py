"""C3: the keyframe loop of examples/semi_dense_vo.py:152-207 on a synthetic sequence with
ground-truth poses -- increment_age -> propagate -> update_depth -> fusion -> regularize per frame,
maps carried from frame to frame -- GPU (device-resident) vs the oracle, bit for bit at every frame.
"""
import numpy as np
import pytest
import torch

import oracle as ora
from tadataka_b200 import scene

pytestmark = pytest.mark.gpu


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_keyframe_loop_matches_oracle_frame_by_frame():
    from tadataka_b200 import device
    H, W, n_frames = 120, 160, 8
    cam = np.array([130.0, 130.0, 79.5, 59.5])
    frames, cam, params = scene.make_sequence(3000, n_frames=n_frames, H=H, W=W, cam=cam)
    oparams = ora.make_params(*params)
    rng = np.random.default_rng(3000)
    default_depth, default_variance, bias = 200.0, 100.0, 1e-6
    # examples/semi_dense_vo.py:210-219: random initial depth, constant variance, age 0
    depth0 = rng.uniform(params[0], params[1], (H, W))
    var0 = np.full((H, W), default_variance)
    age0 = np.zeros((H, W), dtype=np.uint64)

    g_depth, g_var, g_age = cu(depth0), cu(var0), cu(age0)
    o_depth, o_var, o_age = depth0.copy(), var0.copy(), age0.copy()
    g_refs, o_refs, poses = [cu(frames[0][0])], [frames[0][0]], [frames[0][1]]
    g_cam = cu(cam)
    updated = 0
    for i in range(1, n_frames):
        img1, T_w1, _ = frames[i]
        T_w0 = poses[-1]
        T10 = np.linalg.inv(T_w1) @ T_w0
        # ---- oracle
        o_age1 = ora.increment_age(o_age, cam, cam, T10, o_depth)
        o_d1, o_v1 = ora.propagate(T10, cam, cam, o_depth, o_var, default_depth, default_variance, bias)
        R = len(o_refs)
        o_nd, o_nv, o_f = ora.update_depth(img1, cam, T_w1, np.stack(o_refs), np.tile(cam, (R, 1)),
                                           np.stack(poses), o_age1, o_d1, o_v1, oparams)
        eps = np.finfo(np.float64).eps
        ok = o_f == 0
        with np.errstate(all="ignore"):
            ia, ib = 1. / (o_d1 + eps), 1. / (o_nd + eps)
            vs = o_v1 + o_nv
            fd = np.where(ok, 1. / ((ia * o_nv + ib * o_v1) / vs + eps), o_nd)
            fv = np.where(ok, (o_v1 * o_nv) / vs, o_nv)
        o_reg = ora.regularize(fd, fv, o_f)
        # ---- GPU, everything stays on the device
        g_T10 = cu(T10)
        g_age1 = device.increment_age(g_age, g_cam, g_cam, g_T10, g_depth)
        g_d1, g_v1 = device.propagate(g_T10, g_cam, g_cam, g_depth, g_var, default_depth,
                                      default_variance, bias)
        g_img1 = cu(img1)
        g_fd, g_fv, g_f = device.update_depth_frames(g_img1, g_cam, cu(T_w1), g_refs,
                                                     cu(np.tile(cam, (R, 1))), cu(np.stack(poses)),
                                                     g_age1, g_d1, g_v1, params, fuse=True,
                                                     device_tables=bool(i % 2))   # both forms of the pointer tables
        g_reg = device.regularize(g_fd, g_fv, g_f)
        # ---- compare
        assert np.array_equal(g_age1.cpu().numpy(), o_age1), f"age differs at frame {i}"
        assert np.array_equal(g_d1.cpu().numpy(), o_d1), f"propagated depth differs at frame {i}"
        assert np.array_equal(g_v1.cpu().numpy(), o_v1), f"propagated variance differs at frame {i}"
        assert np.array_equal(g_f.cpu().numpy(), o_f), f"flags differ at frame {i}"
        assert np.array_equal(g_fd.cpu().numpy(), fd), f"fused depth differs at frame {i}"
        assert np.array_equal(g_fv.cpu().numpy(), fv), f"fused variance differs at frame {i}"
        assert np.array_equal(g_reg.cpu().numpy(), o_reg), f"regularised depth differs at frame {i}"
        updated += int(ok.sum())
        # carry on (examples/semi_dense_vo.py:203-206)
        g_depth, g_var, g_age = g_reg, g_fv, g_age1
        o_depth, o_var, o_age = o_reg, fv, o_age1
        g_refs.append(g_img1)
        o_refs.append(img1)
        poses.append(T_w1)
    assert updated > 100, "the sequence never produced a successful update"
    assert int(o_age.max()) >= 3


def test_mapper_runs_the_same_loop():
    """tadataka_b200.mapper.SemiDenseMapper == the hand-written loop above == the oracle."""
    from rust_bindings.camera import CameraParameters
    from tadataka_b200.mapper import SemiDenseMapper
    H, W, n_frames = 96, 128, 6
    cam = np.array([100.0, 100.0, 63.5, 47.5])
    frames, cam, params = scene.make_sequence(3100, n_frames=n_frames, H=H, W=W, cam=cam)
    oparams = ora.make_params(*params)
    cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
    rng = np.random.default_rng(3100)
    depth0 = rng.uniform(params[0], params[1], (H, W))
    m = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0,
                        fuse=False, regularize=False)
    o_depth, o_var, o_age = depth0.copy(), np.full((H, W), 100.0), np.zeros((H, W), dtype=np.uint64)
    refs, poses = [frames[0][0]], [frames[0][1]]
    for i in range(1, n_frames):
        img1, T_w1, _ = frames[i]
        T10 = np.linalg.inv(T_w1) @ poses[-1]
        o_age = ora.increment_age(o_age, cam, cam, T10, o_depth)
        d1, v1 = ora.propagate(T10, cam, cam, o_depth, o_var, 200.0, 100.0, 1e-6)
        R = len(refs)
        o_depth, o_var, o_f = ora.update_depth(img1, cam, T_w1, np.stack(refs), np.tile(cam, (R, 1)),
                                               np.stack(poses), o_age, d1, v1, oparams)
        refs.append(img1)
        poses.append(T_w1)
        f = m.track(cp, img1, T_w1)
        depth, var, age, flag = m.maps()
        assert np.array_equal(flag, o_f) and np.array_equal(f.cpu().numpy(), o_f)
        assert np.array_equal(depth, o_depth) and np.array_equal(var, o_var)
        assert np.array_equal(age, o_age)
    assert m.n_frames == n_frames


def test_pruning_unused_reference_frames_changes_nothing_but_the_memory():
    """SemiDenseMapper(prune=True) drops the reference frames no pixel's age points to any more
    (the reference's TODO, examples/semi_dense_vo.py:202): same maps as the mapper that keeps
    every frame, fewer frames held once old pixels have left the map."""
    from rust_bindings.camera import CameraParameters
    from tadataka_b200 import device
    from tadataka_b200.mapper import SemiDenseMapper
    H, W, n_frames = 96, 128, 9
    cam = np.array([100.0, 100.0, 63.5, 47.5])
    frames, cam, params = scene.make_sequence(3200, n_frames=n_frames, H=H, W=W, cam=cam)
    cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
    depth0 = np.random.default_rng(3200).uniform(params[0], params[1], (H, W))
    keep = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0)
    lean = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0,
                           prune=True)
    for i in range(1, n_frames):
        img1, T_w1, _ = frames[i]
        keep.track(cp, img1, T_w1)
        lean.track(cp, img1, T_w1)
        for a, b in zip(keep.maps(), lean.maps()):
            assert np.array_equal(a, b, equal_nan=True)
        counts = device.age_usage(keep.age, i).cpu().numpy()
        assert counts.sum() == H * W and counts[-1] == 0
        assert np.array_equal(counts[:-1], np.bincount(keep.maps()[2].ravel().astype(np.int64), minlength=i + 1))
        if i == 4:   # a re-initialisation: every pixel young again, in both mappers
            for m in (keep, lean):
                m.age = torch.clamp(m.age.to(torch.int64), max=1).to(torch.uint64)
            lean._drop_unused_frames()
            assert lean.n_frames == 2 and keep.n_frames == 5
    assert lean.n_frames < keep.n_frames == n_frames


def test_8bit_frames_are_converted_on_the_device_bit_for_bit():
    """A mapper fed 8-bit gray / RGB / RGBA frames (converted by ingest_u8_kernel) holds exactly the
    f64 images the host conversion gives -- v / 255 and the rgb2gray luma the reference's loaders
    apply (tadataka/dataset/new_tsukuba.py:85-86), evaluated left to right -- and produces the same
    maps as a mapper fed those f64 images."""
    from rust_bindings.camera import CameraParameters
    from tadataka_b200.mapper import SemiDenseMapper
    H, W, n_frames = 96, 128, 5
    cam = np.array([100.0, 100.0, 63.5, 47.5])
    frames, cam, params = scene.make_sequence(3400, n_frames=n_frames, H=H, W=W, cam=cam)
    cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
    rng = np.random.default_rng(3400)
    depth0 = rng.uniform(params[0], params[1], (H, W))
    gray8 = [np.round(f[0] * 255.0).astype(np.uint8) for f in frames]
    rgba8 = []
    for g in gray8:   # a colour frame whose channels differ, plus an alpha channel that must be ignored
        c = np.stack([g, np.roll(g, 1, axis=1), 255 - g, rng.integers(0, 256, g.shape).astype(np.uint8)], axis=-1)
        rgba8.append(np.ascontiguousarray(c))

    def luma(c):
        r, g, b = (c[..., k].astype(np.float64) / 255.0 for k in range(3))
        return (r * 0.2125 + g * 0.7154) + b * 0.0721

    for kind, raw, want in (("gray", gray8, [g.astype(np.float64) / 255.0 for g in gray8]),
                            ("rgba", rgba8, [luma(c) for c in rgba8]),
                            ("rgb", [np.ascontiguousarray(c[..., :3]) for c in rgba8], [luma(c) for c in rgba8])):
        a = SemiDenseMapper(cp, raw[0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0, fuse=True)
        b = SemiDenseMapper(cp, want[0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0, fuse=True)
        for i in range(1, n_frames):
            a.track(cp, raw[i], frames[i][1])
            b.track(cp, want[i], frames[i][1])
        a.sync()
        held = a.reference_images().cpu().numpy()
        assert held.shape == (n_frames, H, W)
        for i in range(n_frames):
            assert np.array_equal(held[i], want[i]), f"{kind} frame {i} differs from the host conversion"
        for x, y in zip(a.maps(), b.maps()):
            assert np.array_equal(x, y, equal_nan=True)
    # the matmul form skimage uses agrees with the left-to-right luma to an ulp or two
    from tests.golden_util import rgb2gray
    assert np.abs(rgb2gray(rgba8[0][..., :3]) - luma(rgba8[0])).max() < 1e-15


def test_mapper_ring_compacts_and_reports_a_full_ring():
    """capacity 4 with pruning: the ring's window reaches the end of the arena and is moved to the
    front without changing the maps; without pruning the fifth frame is refused."""
    from rust_bindings.camera import CameraParameters
    from tadataka_b200.mapper import SemiDenseMapper
    H, W, n_frames = 64, 96, 10
    cam = np.array([80.0, 80.0, 47.5, 31.5])
    frames, cam, params = scene.make_sequence(3500, n_frames=n_frames, H=H, W=W, cam=cam)
    cp = CameraParameters((cam[0], cam[1]), (cam[2], cam[3]))
    depth0 = np.random.default_rng(3500).uniform(params[0], params[1], (H, W))
    big = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0, capacity=32)
    small = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0, prune=True,
                            capacity=4)
    for i in range(1, n_frames):
        big.track(cp, frames[i][0], frames[i][1])
        small.track(cp, frames[i][0], frames[i][1])
        if i % 2 == 0:   # keep the map young so that old frames can go
            for m in (big, small):
                m.age = torch.clamp(m.age.to(torch.int64), max=1).to(torch.uint64)
            small._drop_unused_frames()
        for x, y in zip(big.maps(), small.maps()):
            assert np.array_equal(x, y, equal_nan=True)
    assert small.n_frames <= 4 and big.n_frames == n_frames
    full = SemiDenseMapper(cp, frames[0][0], frames[0][1], params, 200.0, 100.0, 1e-6, depth_map=depth0, capacity=4)
    for i in range(1, 4):
        full.track(cp, frames[i][0], frames[i][1])
    with pytest.raises(Exception, match="larger capacity"):
        full.track(cp, frames[4][0], frames[4][1])
