"""GPU parity: update_depth through the C-ABI vs the oracle on the same seeded inputs.

Gates (BASELINE.json north_star / SURVEY 8d): flag map and search argmin agree on >= 99.9 % of
pixels; inverse depth and variance within 1e-4 relative where both succeed.  The kernels are
written to evaluate every expression in the oracle's order without FMA, so the tests first assert
the much stronger bit-exact equality and print the gate figures.
"""
import numpy as np
import pytest
import torch

import oracle as ora
from tadataka_b200 import scene

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4        # north star: inverse depth and variance within 1e-4 relative
AGREE_MIN = 0.999     # north star: flags / argmin agree on >= 99.9 % of pixels


def _dev(b):
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in b.items()
            if k not in ("params", "gt_depth")}


def _oracle(b, threads=8):
    return ora.update_depth_batch(b["key_img"], b["key_cam"], b["T_wk"], b["ref_imgs"],
                                  b["ref_cams"], b["T_wr"], b["age"], b["prior_depth"],
                                  b["prior_variance"], ora.make_params(*b["params"]),
                                  threads=threads, debug=True)


def _gpu(b, **kw):
    from tadataka_b200 import device
    t = _dev(b)
    r = device.update_depth(t["key_img"], t["key_cam"], t["T_wk"], t["ref_imgs"], t["ref_cams"],
                            t["T_wr"], t["age"], t["prior_depth"], t["prior_variance"], b["params"],
                            debug=True, **kw)
    torch.cuda.synchronize()
    return [x.cpu().numpy() for x in r]


def assert_parity(gpu, orc, exact=True):
    d, v, f, am = gpu[:4]
    od, ov, of, oam = orc[:4]
    flag_agree = float((f == of).mean())
    both = (f == 0) & (of == 0)
    arg_agree = float((am[both] == oam[both]).mean()) if both.any() else 1.0
    assert flag_agree >= AGREE_MIN, f"flags agree on {flag_agree:.5f}"
    assert arg_agree >= AGREE_MIN, f"argmin agrees on {arg_agree:.5f}"
    m = both & (am == oam)
    inv_g, inv_o = 1.0 / d[m], 1.0 / od[m]
    assert np.all(np.abs(inv_g - inv_o) <= REL_TOL * np.abs(inv_o))
    assert np.all(np.abs(v[m] - ov[m]) <= REL_TOL * np.abs(ov[m]))
    # failed pixels carry the prior through unchanged
    nm = (f != 0) & (of != 0) & (f == of)
    assert np.array_equal(d[nm], od[nm]) and np.array_equal(v[nm], ov[nm])
    if exact:
        assert np.array_equal(f, of), "flag maps not identical"
        assert np.array_equal(am, oam), "argmin maps not identical"
        assert np.array_equal(d, od), "depth maps not bit-identical"
        assert np.array_equal(v, ov), "variance maps not bit-identical"
    return flag_agree, arg_agree


def test_c1_single_pair_vga():
    b = scene.stack_keyframes([scene.make_pair(1000)])
    gpu, orc = _gpu(b), _oracle(b)
    assert (orc[2] == 0).sum() > 10000
    assert_parity(gpu, orc)


@pytest.mark.parametrize("plan", [0, 1, 2])
def test_every_kernel_plan_gives_the_oracle_result(plan):
    """Sieve + length-sorted line records (2, the default), sieve + per-CTA search over a candidate
    queue (0), and one kernel with in-CTA compaction (1)."""
    from tadataka_b200 import _lib
    sets = [scene.stack_keyframes([scene.make_pair(1001, H=240, W=320, cam=scene.CAM_VGA / 2)]),
            scene.stack_keyframes([scene.make_keyframe(1002 + i, R=3, H=101, W=147,
                                                       cam=np.array([120., 118., 73., 50.])) for i in range(3)]),
            scene.stack_keyframes([scene.make_pair(4100, H=360, W=640, cam=scene.CAM_720P / 2, pset=scene.SET_T)])]
    try:
        assert _lib.set_kernel_plan(plan) == 0
        for b in sets:
            gpu = _gpu(b, hist=True)
            orc = _oracle(b)
            assert_parity(gpu, orc)
            assert [int(h) for h in gpu[4].sum(0)] == [int((orc[2] == -k).sum()) for k in range(10)]
            fused = _gpu(b, fuse=True)
            want_d, want_v = _fuse_reference(b, orc)
            assert np.array_equal(fused[0], want_d) and np.array_equal(fused[1], want_v)
    finally:
        _lib.set_kernel_plan(_lib.DEFAULT_PLAN)


def test_c2_keyframe_with_8_references_by_age():
    b = scene.stack_keyframes([scene.make_keyframe(2000, R=8)])
    gpu, orc = _gpu(b, hist=True), _oracle(b)
    assert (orc[2] == -9).sum() > 0 and (orc[2] == 0).sum() > 1000
    assert_parity(gpu, orc)
    hist = gpu[4][0]
    assert [int(h) for h in hist] == [int((orc[2] == -k).sum()) for k in range(10)]


def test_c4_720p_tum_scale_long_search_lines():
    b = scene.stack_keyframes([scene.make_pair(4000 + i, H=720, W=1280, cam=scene.CAM_720P,
                                               pset=scene.SET_T) for i in range(2)])
    gpu, orc = _gpu(b), _oracle(b)
    assert orc[4].max() > 40          # search lines longer than one 32-lane chunk
    assert_parity(gpu, orc)


def test_batch_of_pairs_matches_one_by_one():
    items = [scene.make_pair(5000 + i, H=120, W=160, cam=scene.CAM_VGA / 4) for i in range(5)]
    b = scene.stack_keyframes(items)
    gpu = _gpu(b)
    assert_parity(gpu, _oracle(b))
    for i, it in enumerate(items):
        one = _gpu(scene.stack_keyframes([it]))
        for a, c in zip(one, gpu):
            assert np.array_equal(a[0], c[i])


def test_ragged_image_size_and_edge_tiles():
    # W, H not multiples of the 64x4 tile
    it = scene.make_keyframe(77, R=3, H=101, W=147, cam=np.array([120., 118., 73., 50.]))
    b = scene.stack_keyframes([it])
    assert_parity(_gpu(b), _oracle(b))


def test_dense_texture_most_pixels_reach_the_search():
    """White-noise images: nearly every pixel passes the gradient test, so the candidate queue holds
    far more than a quarter of the pixels and the search kernel's CTAs loop over several chunks."""
    rng = np.random.default_rng(42)
    it = scene.make_keyframe(43, R=2, H=120, W=192, cam=np.array([150., 150., 95.5, 59.5]))
    it["key_img"] = rng.random((120, 192))
    it["ref_imgs"] = rng.random((2, 120, 192))
    b = scene.stack_keyframes([it, it])
    gpu, orc = _gpu(b), _oracle(b)
    assert (orc[3] >= 0).mean() > 0.5          # more than half of all pixels were searched
    assert_parity(gpu, orc)


def test_general_poses_distinct_cameras_and_tiny_images():
    """Key pose != identity, every reference with its own intrinsics, images smaller than a tile."""
    rng = np.random.default_rng(77)
    it = scene.make_keyframe(78, R=3, H=90, W=130, cam=np.array([115., 112., 64.2, 44.7]))
    # move the whole rig by a rigid transform G: geometry is unchanged, the numbers are not
    G = scene.pose(scene.rotation(0.3, -0.7, 1.1), (12.0, -40.0, 7.5))
    it["T_wk"] = G @ it["T_wk"]
    it["T_wr"] = np.stack([G @ T for T in it["T_wr"]])
    it["ref_cams"] = it["ref_cams"] * rng.uniform(0.97, 1.03, it["ref_cams"].shape)
    b = scene.stack_keyframes([it])
    orc = _oracle(b)
    assert (orc[2] == 0).sum() > 100
    assert_parity(_gpu(b), orc)
    for H, W in [(1, 1), (3, 5), (5, 70), (9, 9)]:
        small = scene.make_keyframe(80 + H, R=2, H=H, W=W, cam=np.array([20., 20., (W - 1) / 2., (H - 1) / 2.]))
        small["key_img"] = rng.random((H, W))
        small["ref_imgs"] = rng.random((2, H, W))
        small["age"] = rng.integers(0, 3, (H, W)).astype(np.uint64)
        bs = scene.stack_keyframes([small, small])
        assert_parity(_gpu(bs), _oracle(bs))


def test_many_small_keyframes_in_one_launch():
    """P = 300 keyframes (grid.z), one reference each."""
    it = scene.make_pair(90, H=24, W=64, cam=np.array([50., 50., 31.5, 11.5]))
    rng = np.random.default_rng(91)
    items = []
    for k in range(300):
        c = dict(it)
        c["prior_variance"] = 10.0 ** rng.uniform(-3.0, 0.0, size=(24, 64))
        items.append(c)
    b = scene.stack_keyframes(items)
    assert_parity(_gpu(b, hist=True), _oracle(b))


def test_every_flag_is_reachable_and_matches():
    """Hand-made priors that drive pixels into each failure branch (hypothesis.rs:21-37,
    semi_dense.rs:27-40,62-81): negative prior, out-of-range hypothesis, behind the reference."""
    it = scene.make_keyframe(11, R=2, H=96, W=128, cam=np.array([100., 100., 63.5, 47.5]))
    d = it["prior_depth"]
    v = it["prior_variance"]
    d[0:8, :] = -5.0                      # -> NEGATIVE_PRIOR_DEPTH
    d[8:16, :] = 5.0                      # inv depth 0.2, far above vmax=1/60
    v[8:16, :] = 1e-4                     # -> HYPOTHESIS_OUT_OF_SEARCH_RANGE
    d[16:24, :] = 0.0                     # inv(0) = 1/eps, huge
    v[24:32, :] = 0.0                     # zero-width search -> REF_EPIPOLAR_TOO_SHORT
    it["age"][32:40, :] = 0               # -> NOT_PROCESSED
    # a reference camera that looks backwards: every point is behind it -> NEGATIVE_REF_DEPTH
    it["T_wr"][0] = scene.pose(scene.rotation(0., np.pi, 0.), (0., 0., 50.))
    it["age"][40:48, :] = 2               # uses reference 0
    b = scene.stack_keyframes([it])
    gpu, orc = _gpu(b), _oracle(b)
    seen = set(np.unique(orc[2]).tolist())
    assert {-9, -8, -7, -6, -5, -1, 0} <= seen, seen
    assert_parity(gpu, orc)


def test_nan_and_inf_inputs_do_not_break_parity():
    it = scene.make_pair(21, H=64, W=128, cam=np.array([100., 100., 63.5, 31.5]))
    it["prior_depth"][3, 5] = np.nan
    it["prior_depth"][4, 6] = np.inf
    it["prior_variance"][5, 7] = np.nan
    it["prior_variance"][6, 8] = np.inf
    it["key_img"][30, 60] = np.nan
    it["ref_imgs"][0][20, 40] = np.nan
    b = scene.stack_keyframes([it])
    gpu, orc = _gpu(b), _oracle(b)
    d, v, f, am = gpu[:4]
    od, ov, of, oam = orc[:4]
    assert np.array_equal(f, of) and np.array_equal(am, oam)
    assert np.array_equal(d, od, equal_nan=True) and np.array_equal(v, ov, equal_nan=True)


def test_age_beyond_reference_count_is_an_error():
    from tadataka_b200 import device
    it = scene.make_keyframe(5, R=2, H=32, W=64, cam=np.array([50., 50., 31.5, 15.5]))
    it["age"][3, 3] = 3
    t = _dev(scene.stack_keyframes([it]))
    with pytest.raises(ValueError, match="Age exceeds"):
        device.update_depth(t["key_img"], t["key_cam"], t["T_wk"], t["ref_imgs"], t["ref_cams"],
                            t["T_wr"], t["age"], t["prior_depth"], t["prior_variance"], it["params"])
    with pytest.raises(ValueError):
        ora.update_depth_batch(*(scene.stack_keyframes([it])[k] for k in
                                 ("key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age",
                                  "prior_depth", "prior_variance")), ora.make_params(*it["params"]))


def test_singular_reference_pose_is_an_error():
    from tadataka_b200 import device
    it = scene.make_pair(6, H=32, W=64, cam=np.array([50., 50., 31.5, 15.5]))
    it["T_wr"][0] = 0.0
    t = _dev(scene.stack_keyframes([it]))
    with pytest.raises(ValueError, match="singular"):
        device.update_depth(t["key_img"], t["key_cam"], t["T_wk"], t["ref_imgs"], t["ref_cams"],
                            t["T_wr"], t["age"], t["prior_depth"], t["prior_variance"], it["params"])


def test_host_pointer_entry_point_matches_device_entry_point():
    """tdk_update_depth_host (what a numpy caller binds), incl. its chunked pipeline."""
    import ctypes as C
    from tadataka_b200 import _lib
    items = [scene.make_keyframe(300 + i, R=3, H=240, W=320, cam=scene.CAM_VGA / 2) for i in range(7)]
    b = scene.stack_keyframes(items)
    P, H, W = b["key_img"].shape
    R = b["ref_imgs"].shape[1]
    od, ov = np.empty((P, H, W)), np.empty((P, H, W))
    of = np.empty((P, H, W), dtype=np.int64)
    am = np.empty((P, H, W), dtype=np.int32)
    hist = np.zeros((P, 10), dtype=np.uint64)

    def p(a):
        return a.ctypes.data_as(C.c_void_p)

    rc = _lib.update_depth_host(p(b["key_img"]), p(b["key_cam"]), p(b["T_wk"]), p(b["ref_imgs"]),
                                p(b["ref_cams"]), p(b["T_wr"]), p(b["age"]), p(b["prior_depth"]),
                                p(b["prior_variance"]), p(b["params"]), P, R, H, W, p(od), p(ov),
                                p(of), p(am), p(hist), 0)
    _lib.check(rc)
    orc = _oracle(b)
    assert_parity([od, ov, of, am], orc)
    assert [int(x) for x in hist.sum(0)] == [int((orc[2] == -k).sum()) for k in range(10)]


def test_host_pointer_entry_point_with_shared_frames():
    """tdk_update_depth_frames_host: frames by pointer, a reference image shared by several
    keyframes uploaded once; 23 keyframes of 6 scenes so that several chunks and streams are used.
    Same maps as the oracle on the equivalent [P,R,H,W] batch."""
    import ctypes as C
    from tadataka_b200 import _lib
    base = [scene.make_keyframe(330 + i, R=3, H=240, W=320, cam=scene.CAM_VGA / 2) for i in range(6)]
    items = []
    rng = np.random.default_rng(33)
    for k in range(23):
        it = dict(base[k % 6])
        it["prior_variance"] = np.ascontiguousarray(10.0 ** rng.uniform(-3.0, 0.0, size=it["prior_variance"].shape))
        items.append(it)
    b = scene.stack_keyframes(items)
    P, H, W = b["key_img"].shape
    R = b["ref_imgs"].shape[1]
    # pointer tables into the six scenes' own images (not into the stacked copy)
    keep = [np.ascontiguousarray(x["key_img"]) for x in base] + [np.ascontiguousarray(x["ref_imgs"]) for x in base]
    kp = (C.c_void_p * P)(*[keep[k % 6].ctypes.data for k in range(P)])
    rp = (C.c_void_p * (P * R))(*[keep[6 + k % 6].ctypes.data + r * H * W * 8 for k in range(P) for r in range(R)])
    od, ov = np.empty((P, H, W)), np.empty((P, H, W))
    of = np.empty((P, H, W), dtype=np.int64)
    am = np.empty((P, H, W), dtype=np.int32)
    hist = np.zeros((P, 10), dtype=np.uint64)

    def p(a):
        return a.ctypes.data_as(C.c_void_p)

    rc = _lib.update_depth_frames_host(kp, p(b["key_cam"]), p(b["T_wk"]), rp, p(b["ref_cams"]), p(b["T_wr"]),
                                       p(b["age"]), p(b["prior_depth"]), p(b["prior_variance"]), p(b["params"]),
                                       P, R, H, W, p(od), p(ov), p(of), p(am), p(hist), 0)
    _lib.check(rc)
    orc = _oracle(b)
    assert_parity([od, ov, of, am], orc)
    assert [int(x) for x in hist.sum(0)] == [int((orc[2] == -k).sum()) for k in range(10)]
    bad = (C.c_void_p * P)(*([None] + [keep[0].ctypes.data] * (P - 1)))
    assert _lib.update_depth_frames_host(bad, p(b["key_cam"]), p(b["T_wk"]), rp, p(b["ref_cams"]), p(b["T_wr"]),
                                         p(b["age"]), p(b["prior_depth"]), p(b["prior_variance"]), p(b["params"]),
                                         P, R, H, W, p(od), p(ov), p(of), None, None, 0) == _lib.ERR_INVALID


def test_frames_host_byte_wire_ring_and_age_errors():
    """tdk_update_depth_frames_host moves ages and flags as bytes through a pinned ring of 64
    chunks: 70 VGA pairs (one keyframe per chunk) make the ring come round.  Same maps as the
    oracle; an age beyond R is still reported, whether it fits a byte or not."""
    import ctypes as C
    from tadataka_b200 import _lib
    base = [scene.make_pair(770 + i) for i in range(2)]
    rng = np.random.default_rng(77)
    items = []
    for k in range(70):
        it = dict(base[k % 2])
        it["prior_variance"] = np.ascontiguousarray(10.0 ** rng.uniform(-3.0, 0.0, size=it["prior_variance"].shape))
        it["age"] = np.ascontiguousarray((rng.random(it["age"].shape) < 0.9).astype(np.uint64))
        items.append(it)
    b = scene.stack_keyframes(items)
    P, H, W = b["key_img"].shape
    R = b["ref_imgs"].shape[1]
    keep = [np.ascontiguousarray(x["key_img"]) for x in base] + [np.ascontiguousarray(x["ref_imgs"]) for x in base]
    kp = (C.c_void_p * P)(*[keep[k % 2].ctypes.data for k in range(P)])
    rp = (C.c_void_p * (P * R))(*[keep[2 + k % 2].ctypes.data + r * H * W * 8 for k in range(P) for r in range(R)])
    od, ov = np.empty((P, H, W)), np.empty((P, H, W))
    of = np.full((P, H, W), 99, dtype=np.int64)
    am = np.empty((P, H, W), dtype=np.int32)

    def p(a):
        return a.ctypes.data_as(C.c_void_p)

    def call(age):
        return _lib.update_depth_frames_host(kp, p(b["key_cam"]), p(b["T_wk"]), rp, p(b["ref_cams"]), p(b["T_wr"]),
                                             p(age), p(b["prior_depth"]), p(b["prior_variance"]), p(b["params"]),
                                             P, R, H, W, p(od), p(ov), p(of), p(am), None, 0)

    _lib.check(call(b["age"]))
    orc = _oracle(b)
    assert_parity([od, ov, of, am], orc)
    for bad_age in (2, 255, 300, 2 ** 40 + 1):     # 2**40 + 1 would read as 1 if the wire just truncated
        age = b["age"].copy()
        age[69, 17, 23] = bad_age
        assert call(age) == _lib.ERR_AGE
    _lib.check(call(b["age"]))                     # and the entry point works again afterwards
    assert_parity([od, ov, of, am], orc)


def test_maps_host_entry_point_pageable_and_pinned():
    """tdk_update_depth_maps_host: frames on the device, maps in host memory.  8 VGA keyframes make
    every map 19.7 MB, so the 16 MB pinned ring of the pageable path comes round; the same call
    with pinned maps takes the direct path.  Both equal the oracle, argmin and histogram included."""
    import ctypes as C
    from tadataka_b200 import _lib
    items = [scene.make_keyframe(880 + i, R=2) for i in range(2)] * 4
    rng = np.random.default_rng(88)
    items = [dict(it, prior_variance=np.ascontiguousarray(10.0 ** rng.uniform(-3.0, 0.0, size=it["prior_variance"].shape)))
             for it in items]
    b = scene.stack_keyframes(items)
    P, H, W = b["key_img"].shape
    R = b["ref_imgs"].shape[1]
    g_key, g_ref = torch.from_numpy(b["key_img"]).cuda(), torch.from_numpy(b["ref_imgs"]).cuda()
    torch.cuda.synchronize()
    kp = (C.c_void_p * P)(*[g_key[k].data_ptr() for k in range(P)])
    rp = (C.c_void_p * (P * R))(*[g_ref[k, r].data_ptr() for k in range(P) for r in range(R)])
    orc = _oracle(b)

    def p(a):
        return C.c_void_p(a.data_ptr()) if isinstance(a, torch.Tensor) else a.ctypes.data_as(C.c_void_p)

    for pinned in (False, True):
        def host(a):
            return torch.from_numpy(a).pin_memory() if pinned else a
        age, pd, pv = host(b["age"].view(np.int64)), host(b["prior_depth"]), host(b["prior_variance"])
        od, ov = host(np.empty((P, H, W))), host(np.empty((P, H, W)))
        of = host(np.full((P, H, W), 99, dtype=np.int64))
        am = host(np.empty((P, H, W), dtype=np.int32))
        hist = np.zeros((P, 10), dtype=np.uint64)
        rc = _lib.update_depth_maps_host(kp, p(b["key_cam"]), p(b["T_wk"]), rp, p(b["ref_cams"]), p(b["T_wr"]), p(age),
                                         p(pd), p(pv), p(b["params"]), P, R, H, W, p(od), p(ov), p(of), p(am), p(hist), 0)
        _lib.check(rc)
        got = [x.numpy() if isinstance(x, torch.Tensor) else x for x in (od, ov, of, am)]
        assert_parity(got, orc)
        assert [int(x) for x in hist.sum(0)] == [int((orc[2] == -k).sum()) for k in range(10)]
    bad = b["age"].copy()
    bad[3, 5, 7] = 2 ** 33
    assert _lib.update_depth_maps_host(kp, p(b["key_cam"]), p(b["T_wk"]), rp, p(b["ref_cams"]), p(b["T_wr"]), p(bad),
                                       p(b["prior_depth"]), p(b["prior_variance"]), p(b["params"]), P, R, H, W,
                                       p(od), p(ov), p(of), None, None, 0) == _lib.ERR_AGE


def _fuse_reference(b, orc):
    """prior (+) oracle measurement where SUCCESS, composed from IEEE numpy ops in the order of
    propagation.rs:27-34 / fusion.rs:3-11."""
    od, ov, of = orc[:3]
    eps = np.finfo(np.float64).eps
    pd, pv = b["prior_depth"], b["prior_variance"]
    with np.errstate(all="ignore"):
        ia, ib = 1. / (pd + eps), 1. / (od + eps)
        vs = pv + ov
        fd = 1. / ((ia * ov + ib * pv) / vs + eps)
        fv = (pv * ov) / vs
    ok = of == 0
    return np.where(ok, fd, od), np.where(ok, fv, ov)


def test_fused_update_equals_update_then_fusion():
    """fuse_on_success=1 (one kernel) == update_depth + tdk_fusion_dev(depth_domain) (two kernels)
    == the oracle's measurement fused with the prior."""
    from tadataka_b200 import device
    b = scene.stack_keyframes([scene.make_keyframe(2100 + i, R=4, H=240, W=320, cam=scene.CAM_VGA / 2)
                               for i in range(3)])
    orc = _oracle(b)
    want_d, want_v = _fuse_reference(b, orc)
    fused = _gpu(b, fuse=True)
    assert np.array_equal(fused[2], orc[2]) and np.array_equal(fused[3], orc[3])
    assert np.array_equal(fused[0], want_d) and np.array_equal(fused[1], want_v)
    t = _dev(b)
    d, v, f = device.update_depth(t["key_img"], t["key_cam"], t["T_wk"], t["ref_imgs"], t["ref_cams"],
                                  t["T_wr"], t["age"], t["prior_depth"], t["prior_variance"], b["params"])
    d2, v2 = device.fusion(t["prior_depth"], d, t["prior_variance"], v, flag=f, depth_domain=True)
    # where not SUCCESS the two-kernel form keeps (mu1, var1) = the prior; update_depth itself stores
    # inv(inv(prior)) for failed estimates, so compare on SUCCESS pixels and on the flags
    ok = orc[2] == 0
    assert np.array_equal(d2.cpu().numpy()[ok], want_d[ok])
    assert np.array_equal(v2.cpu().numpy()[ok], want_v[ok])
    # fusing shrinks the variance
    assert np.all(want_v[ok] <= b["prior_variance"][ok])


def test_full_size_c2_batch_properties():
    """BASELINE.json configs[1] at the bench's full size (64 VGA keyframes x 8 references): too big
    for the oracle in a test, so size-independent properties instead -- the three kernel plans (three
    different sets of kernels) agree bit for bit, a rerun is bit-identical, the flag histogram adds up to
    the pixel count, failed pixels carry their prior -- plus the oracle on three keyframes of the
    batch."""
    import bench
    from tadataka_b200 import _lib
    b = bench.make_workload(0, 64, 8)
    P, H, W = b["key_img"].shape
    assert (P, H, W) == (64, 480, 640)
    try:
        _lib.set_kernel_plan(2)
        a0 = _gpu(b, hist=True)
        a1 = _gpu(b, hist=True)
        _lib.set_kernel_plan(0)
        b0 = _gpu(b, hist=True)
        _lib.set_kernel_plan(1)
        c0 = _gpu(b, hist=True)
    finally:
        _lib.set_kernel_plan(_lib.DEFAULT_PLAN)
    for x, y, z, w in zip(a0, a1, b0, c0):
        assert np.array_equal(x, y, equal_nan=True), "rerun differs"
        assert np.array_equal(x, z, equal_nan=True), "kernel plans 2 and 0 differ"
        assert np.array_equal(x, w, equal_nan=True), "kernel plans 2 and 1 differ"
    d, v, f, am, hist = a0
    assert f.min() >= -9 and f.max() <= 0
    assert int(hist.sum()) == P * H * W
    assert [int(h) for h in hist.sum(0)] == [int((f == -k).sum()) for k in range(10)]
    failed = f != 0
    assert np.array_equal(v[failed], b["prior_variance"][failed])
    untouched = (f == -9) | (f == -7) | (f == -1)
    assert np.array_equal(d[untouched], b["prior_depth"][untouched])
    assert (am[f == 0] >= 2).all() and (am[(f != 0) & (f != -7) & (f != -1)] <= am.max()).all()
    assert (f == 0).mean() > 0.05
    sub = [0, 13, 63]
    bs = {k: (np.ascontiguousarray(val[sub]) if k != "params" else val) for k, val in b.items()}
    orc = _oracle(bs)
    for g, o in zip((d[sub], v[sub], f[sub], am[sub]), orc[:4]):
        assert np.array_equal(g, o, equal_nan=True)


def test_newtsukuba_golden_pair():
    """The reference's own fixture (tests/vo/semi_dense/test_semi_dense.py:40-73) with the values
    the oracle pinned in tests/golden/."""
    from tests.golden_util import load_newtsukuba
    g = load_newtsukuba()
    H, W = g["key"].shape
    b = dict(key_img=g["key"][None], key_cam=g["cam"][None], T_wk=g["T_wl"][None],
             ref_imgs=g["ref"][None, None], ref_cams=g["cam"][None, None], T_wr=g["T_wr"][None, None],
             age=np.ones((1, H, W), dtype=np.uint64), prior_depth=np.full((1, H, W), 200.0),
             prior_variance=np.ones((1, H, W)), params=np.array([60.0, 1000.0, 0.01, 0.01, 0.01, 0.2]))
    d, v, f, am = _gpu(b)[:4]
    assert [int((f == -k).sum()) for k in range(9, -1, -1)] == g["flag_hist"].tolist()
    assert np.array_equal(f[0], g["flag_map"].astype(np.int64))
    assert np.array_equal(am[0], g["argmin_map"].astype(np.int32))
    ok = f[0] == 0
    assert np.array_equal(d[0][ok], g["depth_success"])
    assert np.array_equal(v[0][ok], g["variance_success"])


def test_shared_denominator_division_is_the_plain_division():
    """tdk::div2 / tdk::div5 (several numerators, one denominator; tdk_common.cuh) must return what
    `/` returns, bit for bit: 3 x 20 M seeded cases over ordinary magnitudes, arbitrary bit patterns
    (NaN, Inf, subnormals) and exponents at the limits of the fast path."""
    import torch
    from tadataka_b200 import _lib
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    for seed in (1, 2, 3):
        _lib.check(_lib.selftest_shared_division_dev(20_000_000, seed, bad.data_ptr(),
                                                     torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        assert int(bad.item()) == 0, f"seed {seed}: {int(bad.item())} quotients differ from a / b"


def test_smallest_accepted_workspace_gives_the_same_maps(monkeypatch):
    """tdk_update_depth_workspace_bytes_min: one keyframe per sub-batch of the sorted plan instead
    of up to 64 -- more launches, identical results; one byte less is refused."""
    import torch
    from tadataka_b200 import _lib, device
    b = scene.stack_keyframes([scene.make_keyframe(2600 + i, R=2, H=96, W=128, cam=scene.CAM_VGA / 5)
                               for i in range(5)])
    try:
        _lib.set_kernel_plan(2)   # the automatic choice would take the two-kernel plan at this size
        want = _gpu(b, hist=True)
        P, H, W = b["key_img"].shape
        small = _lib.update_depth_workspace_bytes_min(P, 2, H, W)
        assert 0 < small < _lib.update_depth_workspace_bytes(P, 2, H, W)
        device._workspaces.clear()
        monkeypatch.setattr(_lib, "update_depth_workspace_bytes", lambda *a: small)
        got = _gpu(b, hist=True)
        for x, y in zip(want, got):
            assert np.array_equal(x, y, equal_nan=True)
        device._workspaces.clear()
        monkeypatch.setattr(_lib, "update_depth_workspace_bytes", lambda *a: small - 1)
        with pytest.raises(_lib.TdkError, match="workspace"):
            _gpu(b)
    finally:
        device._workspaces.clear()
        _lib.set_kernel_plan(_lib.DEFAULT_PLAN)


@pytest.mark.parametrize("min_gradient", [0.0, -1.0, 1e-200, 1e-3, 0.05, 0.2, 0.7, 1.5, 1e200, float("inf"),
                                          float("nan"), 0.2 * (1 + 2 ** -52), 0.2 * (1 - 2 ** -53)])
def test_gradient_threshold_without_square_roots_matches_the_oracle(min_gradient):
    """The sieve tests s < T instead of sqrt(s) < min_gradient, T found on the host: every kind of
    threshold (non-positive, tiny, ordinary, neighbours of an ordinary one, huge, inf, NaN) must
    give the oracle's flags and maps bit for bit, in every plan."""
    from tadataka_b200 import _lib
    b = scene.stack_keyframes([scene.make_keyframe(2700, R=2, H=96, W=128, cam=scene.CAM_VGA / 5)])
    b["params"] = b["params"].copy()
    b["params"][5] = min_gradient
    orc = _oracle(b)
    try:
        for plan in (2, 0, 1):
            _lib.set_kernel_plan(plan)
            gpu = _gpu(b)
            for x, y in zip(gpu[:4], orc[:4]):
                assert np.array_equal(x, y, equal_nan=True), f"plan {plan}"
    finally:
        _lib.set_kernel_plan(_lib.DEFAULT_PLAN)


def test_update_depth_replays_from_a_cuda_graph():
    """The launch is memset nodes and kernels on the caller's stream only, so it can be captured once
    and replayed: the replay must write the maps an eager call writes, also after the inputs changed
    in place."""
    import torch
    from tadataka_b200 import device
    items = [scene.make_keyframe(2800 + i, R=3, H=96, W=128, cam=scene.CAM_VGA / 5) for i in range(3)]
    b = scene.stack_keyframes(items)
    t = _dev(b)
    args = [t[k] for k in ("key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth",
                           "prior_variance")]
    P, H, W = b["key_img"].shape
    out = (torch.empty((P, H, W), dtype=torch.float64, device="cuda"),
           torch.empty((P, H, W), dtype=torch.float64, device="cuda"),
           torch.empty((P, H, W), dtype=torch.int64, device="cuda"))
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        device.update_depth(*args, b["params"], check=False, out=out, fuse=True)   # allocates the workspace
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            device.update_depth(*args, b["params"], check=False, out=out, fuse=True)
    for trial in range(2):
        for o in out:
            o.zero_()
        g.replay()
        torch.cuda.synchronize()
        want = device.update_depth(*args, b["params"], fuse=True)
        torch.cuda.synchronize()
        for x, y in zip(out, want):
            assert torch.equal(x, y) or np.array_equal(x.cpu().numpy(), y.cpu().numpy(), equal_nan=True)
        assert int((out[2] == 0).sum()) > 20
        # new priors in the same buffers: the graph must pick them up
        t["prior_variance"].mul_(0.5)


def test_subpixel_refinement_is_off_by_default_and_lowers_the_depth_error():
    """north star step 3 (opt-in; the reference has none).  Default: nothing changes.
    Switched on: flags and search indices stay, depths of the successful pixels move, and against
    the synthetic scene's exact depth their median error shrinks in every plan."""
    from tadataka_b200 import _lib, device
    it = scene.make_pair(1234, H=240, W=320, cam=np.array([260.0, 259.0, 159.5, 119.5]))
    b = scene.stack_keyframes([it, scene.make_pair(1235, H=240, W=320, cam=np.array([260.0, 259.0, 159.5, 119.5]))])
    names = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth", "prior_variance"]
    t = _dev(b)
    gt = b["gt_depth"]
    try:
        for plan in (2, 0, 1):
            _lib.check(_lib.set_kernel_plan(plan))
            d0, v0, f0, a0 = device.update_depth(*(t[k] for k in names), b["params"], debug=True)
            _lib.check(_lib.set_subpixel_refinement(1))
            d1, v1, f1, a1 = device.update_depth(*(t[k] for k in names), b["params"], debug=True)
            _lib.check(_lib.set_subpixel_refinement(0))
            d2, v2, f2, a2 = device.update_depth(*(t[k] for k in names), b["params"], debug=True)
            assert torch.equal(d0, d2) and torch.equal(v0, v2) and torch.equal(f0, f2) and torch.equal(a0, a2)
            d0, d1, f0, f1 = d0.cpu().numpy(), d1.cpu().numpy(), f0.cpu().numpy(), f1.cpu().numpy()
            assert np.array_equal(a0.cpu().numpy(), a1.cpu().numpy())
            both = (f0 == 0) & (f1 == 0)
            assert both.sum() > 1000 and (f0 != f1).mean() < 1e-3
            assert (d0[both] != d1[both]).mean() > 0.5
            e0, e1 = np.abs(d0[both] - gt[both]), np.abs(d1[both] - gt[both])
            assert np.median(e1) < np.median(e0), (plan, np.median(e0), np.median(e1))
            assert e1.mean() < e0.mean(), (plan, e0.mean(), e1.mean())
    finally:
        _lib.set_subpixel_refinement(0)
        _lib.set_kernel_plan(_lib.DEFAULT_PLAN)
