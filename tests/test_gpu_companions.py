"""GPU parity of the companion kernels and the drop-in helper surface vs the oracle."""
import numpy as np
import pytest
import torch

import oracle as ora
from tadataka_b200 import scene

pytestmark = pytest.mark.gpu


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_fusion_reference_vector():
    # src/semi_dense/fusion.rs test_fusion
    from rust_bindings.semi_dense import fusion
    mu1 = np.array([[1.9, -2.2], [-3.8, 4.1], [-1.5, 4.5]])
    mu2 = np.array([[-4.1, -2.5], [1.2, 5.0], [6.4, 4.1]])
    v1 = np.array([[4.8, 2.2], [3.1, 6.8], [4.0, 2.1]])
    v2 = np.array([[4.2, 3.1], [0.01, 2.0], [6.0, 3.9]])
    mu0, v0 = fusion(mu1, mu2, v1, v2)
    assert np.array_equal(mu0, (v2 * mu1 + v1 * mu2) / (v1 + v2))
    assert np.array_equal(v0, (v1 * v2) / (v1 + v2))


def test_fusion_random_and_gated():
    from tadataka_b200 import device
    rng = np.random.default_rng(0)
    n = (7, 333)
    mu1, mu2 = rng.normal(size=n), rng.normal(size=n)
    v1, v2 = rng.uniform(0.01, 5, n), rng.uniform(0.01, 5, n)
    flag = rng.integers(-9, 1, size=n).astype(np.int64)
    emu, ev = ora.fusion_arrays(mu1, mu2, v1, v2)
    mu, v = device.fusion(cu(mu1), cu(mu2), cu(v1), cu(v2))
    assert np.array_equal(mu.cpu().numpy(), emu) and np.array_equal(v.cpu().numpy(), ev)
    mu, v = device.fusion(cu(mu1), cu(mu2), cu(v1), cu(v2), flag=cu(flag))
    assert np.array_equal(mu.cpu().numpy(), np.where(flag == 0, emu, mu1))
    assert np.array_equal(v.cpu().numpy(), np.where(flag == 0, ev, v1))


def test_regularize_reference_vector_and_random():
    from rust_bindings.semi_dense import regularize
    depth = np.array([[1., 2., 4., 2.], [3., 4., 1., 9.], [1., 4., 8., 1.]])
    var = np.array([[1., 4., 3., 5.], [3., 5., 2., 1.], [2., 4., 2., 2.]])
    flag = np.array([[1, 0, 1, 1], [1, 1, 0, 1], [0, 1, 0, 1]], dtype=np.int64)
    assert np.array_equal(regularize(depth, var, flag), ora.regularize(depth, var, flag))
    rng = np.random.default_rng(1)
    for shape in [(37, 91), (4, 64), (130, 7)]:
        depth, var = rng.uniform(0.5, 50, shape), rng.uniform(0.01, 3, shape)
        flag = np.where(rng.random(shape) < 0.4, 0, rng.integers(-9, 0, shape)).astype(np.int64)
        flag[:3, :5] = -6     # a patch with no Success neighbour keeps its depth
        out = regularize(depth, var, flag)
        assert np.array_equal(out, ora.regularize(depth, var, flag))
        assert out[0, 0] == depth[0, 0]


def test_increment_age_reference_vector():
    # src/semi_dense/age.rs test_increment_age
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import increment_age
    W, H = 12, 16
    cp = CameraParameters((10., 10.), (W / 2., H / 2.))
    T10 = np.eye(4)
    T10[2, 3] = 10.
    age1 = increment_age(np.zeros((H, W), dtype=np.uint64), cp, cp, T10, 10.0 * np.ones((H, W)))
    expected = np.zeros((H, W), dtype=np.uint64)
    expected[4:12, 3:9] = 1
    assert age1.dtype == np.uint64 and np.array_equal(age1, expected)


def test_propagate_reference_vector():
    # src/semi_dense/propagation.rs test_propagate : 8x8 zoom-out, 16 -> 1 collisions
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import propagate
    W = H = 8
    cp = CameraParameters((100., 100.), (W / 2., H / 2.))
    T10 = np.eye(4)
    T10[2, 3] = 300.
    d1, v1 = propagate(T10, cp, cp, np.full((H, W), 100.), np.full((H, W), 20.), 60., 8., 3.)
    exp_d = np.full((H, W), 60.)
    exp_d[3:5, 3:5] = 400.
    np.testing.assert_allclose(d1, exp_d, rtol=0, atol=1e-4)
    exp_v = np.full((H, W), 8.)
    exp_v[3:5, 3:5] = ora.propagate_variance(100., 400., 20., 3.) / 16.
    np.testing.assert_allclose(v1, exp_v, rtol=0, atol=1e-4)
    cam = ora.cam((100., 100.), (W / 2., H / 2.))
    od, ov = ora.propagate(T10, cam, cam, np.full((H, W), 100.), np.full((H, W), 20.), 60., 8., 3.)
    assert np.array_equal(d1, od) and np.array_equal(v1, ov)


def _moving_frames(seed, H, W, cam, zoom):
    it = scene.make_pair(seed, H=H, W=W, cam=cam)
    rng = np.random.default_rng(seed)
    T10 = scene.pose(scene.rotation(0.01, -0.02, 0.005), (3.0, -1.0, zoom))
    depth0 = it["gt_depth"] * (1 + 0.05 * rng.standard_normal((H, W)))
    var0 = rng.uniform(1e-6, 1e-4, (H, W))
    age0 = rng.integers(0, 6, (H, W)).astype(np.uint64)
    return T10, depth0, var0, age0


@pytest.mark.parametrize("zoom", [-40.0, 0.0, 150.0])
def test_propagate_and_age_on_a_moving_camera(zoom):
    """Forward / sideways / backward motion: backward motion (zoom-out) makes many sources collide
    on one target, which exercises the ordered fold and the last-writer rule."""
    from tadataka_b200 import device
    H, W = 120, 160
    cam = np.array([130.0, 130.0, 79.5, 59.5])
    T10, depth0, var0, age0 = _moving_frames(31, H, W, cam, zoom)
    od, ov = ora.propagate(T10, cam, cam, depth0, var0, 250.0, 10.0, 1e-5)
    d1, v1 = device.propagate(cu(T10), cu(cam), cu(cam), cu(depth0), cu(var0), 250.0, 10.0, 1e-5)
    assert np.array_equal(d1.cpu().numpy(), od) and np.array_equal(v1.cpu().numpy(), ov)
    oa = ora.increment_age(age0, cam, cam, T10, depth0)
    a1 = device.increment_age(cu(age0), cu(cam), cu(cam), cu(T10), cu(depth0))
    assert np.array_equal(a1.cpu().numpy(), oa)
    if zoom > 0:
        assert (oa == 0).sum() > 0.2 * H * W      # zoom-out leaves a border unreached


def test_propagate_heavy_collisions_use_the_heap_sort_path():
    """Strong zoom-out: > 16 sources per target."""
    from tadataka_b200 import device
    H, W = 96, 128
    cam = np.array([100.0, 100.0, 63.5, 47.5])
    rng = np.random.default_rng(5)
    depth0 = 100.0 * (1 + 0.3 * rng.random((H, W)))
    var0 = rng.uniform(1e-7, 1e-3, (H, W))
    T10 = scene.pose(None, (0.3, 0.2, 700.0))
    od, ov = ora.propagate(T10, cam, cam, depth0, var0, 60.0, 8.0, 1e-6)
    d1, v1 = device.propagate(cu(T10), cu(cam), cu(cam), cu(depth0), cu(var0), 60.0, 8.0, 1e-6)
    assert np.array_equal(d1.cpu().numpy(), od) and np.array_equal(v1.cpu().numpy(), ov)
    assert (od != 60.0).sum() < H * W / 16


def test_propagate_and_age_batched():
    from tadataka_b200 import device
    H, W, P = 60, 80, 3
    cam = np.array([70.0, 70.0, 39.5, 29.5])
    sets = [_moving_frames(40 + i, H, W, cam, z) for i, z in enumerate([-20., 30., 90.])]
    T10 = np.stack([s[0] for s in sets])
    d0, v0 = np.stack([s[1] for s in sets]), np.stack([s[2] for s in sets])
    a0 = np.stack([s[3] for s in sets])
    cams = np.tile(cam, (P, 1))
    d1, v1 = device.propagate(cu(T10), cu(cams), cu(cams), cu(d0), cu(v0), 250.0, 10.0, 1e-5)
    a1 = device.increment_age(cu(a0), cu(cams), cu(cams), cu(T10), cu(d0))
    for i in range(P):
        od, ov = ora.propagate(T10[i], cam, cam, d0[i], v0[i], 250.0, 10.0, 1e-5)
        assert np.array_equal(d1[i].cpu().numpy(), od) and np.array_equal(v1[i].cpu().numpy(), ov)
        assert np.array_equal(a1[i].cpu().numpy(), ora.increment_age(a0[i], cam, cam, T10[i], d0[i]))
    # the fused entry point (one warp of the map for both) gives the same three maps
    fd, fv, fa = device.propagate_with_age(cu(T10), cu(cams), cu(cams), cu(d0), cu(v0), cu(a0), 250.0, 10.0, 1e-5)
    assert torch.equal(fd, d1) and torch.equal(fv, v1) and torch.equal(fa, a1)


# ------------------------------------------------------------------ helper surface
def test_helpers_match_reference_vectors_and_oracle():
    import rust_bindings.interpolation as ri
    import rust_bindings.projection as rp
    import rust_bindings.transform as rt
    import rust_bindings.triangulation as rtri
    import rust_bindings.warp as rw
    import rust_bindings.homogeneous as rh
    # tests/test_warp.py:77-107 / src/warp.rs test_warp_2d
    T = np.array([[0., 0., 1., 0.], [0., 1., 0., 0.], [-1., 0., 0., 4.], [0., 0., 0., 1.]])
    xs1, d1 = rw.warp_vecs(T, np.array([[0., 0.], [2., -1.]]), np.array([2., 4.]))
    assert np.array_equal(xs1, [[0.5, 0.0], [-1.0, 1.0]]) and np.array_equal(d1, [4., -4.])
    x1, d = rw.warp_vec(T, np.array([0., 0.]), 2.)
    assert np.array_equal(x1, [0.5, 0.]) and d == 4. and isinstance(d, float)
    # src/projection.rs tests
    assert np.array_equal(rp.project_vecs(np.array([[0., 0., 0.], [1., 4., 2.], [-1., 3., 5.]])),
                          [[0., 0.], [0.5, 2.], [-0.2, 0.6]])
    assert np.array_equal(rp.project_vec(np.array([3., 5., 5.])), [0.6, 1.])
    assert np.array_equal(rp.inv_project_vecs(np.array([[0.5, 2.], [-0.2, 0.6]]), np.array([2., 5.])),
                          [[1., 4., 2.], [-1., 3., 5.]])
    assert np.array_equal(rp.inv_project_vec(np.array([0.5, 2.]), 2.), [1., 4., 2.])
    # src/transform.rs test_transform_points
    T2 = np.array([[1., 0, 0, 1], [0, 0, -1, 2], [0, 1, 0, 3], [0, 0, 0, 1]])
    assert np.array_equal(rt.transform(T2, np.array([[1., 2., 5.], [4., -2., 3.]])),
                          [[2., -3., 5.], [5., -1., 1.]])
    # homogeneous.rs
    assert np.array_equal(rh.to_homogeneous_vec(np.array([1., 2.])), [1., 2., 1.])
    assert np.array_equal(rh.from_homogeneous_vecs(np.array([[1., 2., 1.], [3., 4., 1.]])),
                          [[1., 2.], [3., 4.]])
    # random vs oracle
    rng = np.random.default_rng(2)
    Tr = scene.pose(scene.rotation(0.1, -0.2, 0.3), (1., -2., 0.5))
    xs, ds = rng.normal(size=(500, 2)), rng.uniform(1, 9, 500)
    exs1, ed1 = ora.warp_vecs(Tr, xs, ds)
    gxs1, gd1 = rw.warp_vecs(Tr, xs, ds)
    assert np.array_equal(gxs1, exs1) and np.array_equal(gd1, ed1)
    got = np.array([rtri.calc_depth0(Tr, xs[i], exs1[i]) for i in range(5)])
    want = np.array([ora.calc_depth0(Tr, xs[i], exs1[i]) for i in range(5)])
    assert np.array_equal(got, want)
    np.testing.assert_allclose(got, ds[:5], rtol=1e-9)     # triangulation recovers the depth
    # tests/test_interpolation.py:12-80
    img = np.array([[0., 1., 5.], [0., 0., 2.], [4., 3., 2.], [5., 6., 1.]])
    cs = np.array([[1.3, 2.6], [0., 0.], [0., 0.1], [0.1, 0.], [2., 2.9], [1.9, 3.], [2., 3.]])
    assert np.array_equal(ri.interpolation(img, cs), ora.interpolate(img, cs))
    big = rng.random((50, 70))
    cs = np.stack([rng.uniform(0, 69, 3000), rng.uniform(0, 49, 3000)], axis=1)
    cs[:300] = np.floor(cs[:300])
    assert np.array_equal(ri.interpolation(big, cs), ora.interpolate(big, cs))
    for bad in ([3.0, 2.01], [3.01, 2.0], [-0.01, 0.0], [0.0, -0.01]):
        with pytest.raises(ValueError):
            ri.interpolation(img, np.array([bad]))


def test_drop_in_update_depth_and_estimate_debug():
    """The numpy API of rust_bindings.semi_dense on the reference's own fixture
    (tests/vo/semi_dense/test_semi_dense.py:40-134)."""
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import Frame, Params, estimate_debug_, update_depth
    from tadataka.vo.semi_dense.flag import ResultFlag as FLAG
    from tests.golden_util import load_newtsukuba
    g = load_newtsukuba()
    cp = CameraParameters((615., 615.), (320., 240.))
    keyframe, refframe = Frame(cp, g["key"], g["T_wl"]), Frame(cp, g["ref"], g["T_wr"])
    shape = g["key"].shape
    params = Params(min_depth=60.0, max_depth=1000.0, geo_coeff=0.01, photo_coeff=0.01,
                    ref_step_size=0.01, min_gradient=0.2)
    depth, variance, flag = update_depth(keyframe, [refframe], np.ones(shape, dtype=np.uint64),
                                         200.0 * np.ones(shape), np.ones(shape), params)
    assert depth.dtype == np.float64 and variance.dtype == np.float64 and flag.dtype == np.int64
    assert np.array_equal(flag, g["flag_map"].astype(np.int64))
    assert np.array_equal(depth[flag == 0], g["depth_success"])
    with pytest.raises(ValueError, match="Age exceeds"):
        update_depth(keyframe, [refframe], 2 * np.ones(shape, dtype=np.uint64),
                     200.0 * np.ones(shape), np.ones(shape), params)
    with pytest.raises(TypeError):
        update_depth(keyframe, [refframe], np.ones(shape, dtype=np.int64),
                     200.0 * np.ones(shape), np.ones(shape), params)

    params = Params(min_depth=0.1, max_depth=1000.0, geo_coeff=0.01, photo_coeff=0.01,
                    ref_step_size=0.01, min_gradient=0.2)
    cases = [((110, 400), -10.0, 10.0, FLAG.NEGATIVE_PRIOR_DEPTH),
             ((110, 400), 0.05, 0.2, FLAG.HYPOTHESIS_OUT_OF_SERCH_RANGE),
             ((390, 100), 2.0, 0.2, FLAG.INSUFFICIENT_GRADIENT),
             ((0, 200), 2.0, 0.2, FLAG.KEY_OUT_OF_RANGE),
             ((116, 400), 2.0, 0.001, FLAG.REF_EPIPOLAR_TOO_SHORT)]
    for u, d, v, want in cases:
        depth, variance, f = estimate_debug_(np.array(u), d, v, keyframe, refframe, params)
        assert f == want and depth == d and variance == v
    # a SUCCESS pixel: same numbers as the oracle's probe
    ys, xs = np.nonzero(g["flag_map"] == 0)
    u = (int(xs[100]), int(ys[100]))
    p6 = ora.make_params(60.0, 1000.0, 0.01, 0.01, 0.01, 0.2)
    want = ora.estimate_debug(u, 200.0, 1.0, g["key"], g["cam"], g["T_wl"], g["ref"], g["cam"],
                              g["T_wr"], p6)
    got = estimate_debug_(np.array(u), 200.0, 1.0, keyframe, refframe,
                          Params(60.0, 1000.0, 0.01, 0.01, 0.01, 0.2))
    assert got == want and got[2] == 0


def test_host_pointer_companions_batched_through_the_c_abi():
    """tdk_increment_age_host / tdk_propagate_host / tdk_fusion_host / tdk_regularize_host with P = 3
    frames in one call, plain host pointers, against the oracle frame by frame."""
    import ctypes as C
    from tadataka_b200 import _lib
    rng = np.random.default_rng(21)
    P, H, W = 3, 60, 84
    cam = np.array([80., 79., 41.5, 29.5])
    cams = np.ascontiguousarray(np.tile(cam, (P, 1)))
    T10 = np.stack([scene.pose(scene.rotation(0.01 * i, -0.02, 0.005), (0.3 * i, -0.1, 0.4 + 0.2 * i)) for i in range(P)])
    depth0 = rng.uniform(4., 9., (P, H, W))
    var0 = rng.uniform(0.01, 0.5, (P, H, W))
    age0 = rng.integers(0, 5, (P, H, W)).astype(np.uint64)

    def p(a):
        return a.ctypes.data_as(C.c_void_p)

    age1 = np.empty((P, H, W), dtype=np.uint64)
    _lib.check(_lib.increment_age_host(p(age0), p(cams), p(cams), p(T10), p(depth0), P, H, W, p(age1)))
    d1, v1 = np.empty((P, H, W)), np.empty((P, H, W))
    _lib.check(_lib.propagate_host(p(T10), p(cams), p(cams), p(depth0), p(var0), P, H, W, 7.0, 50.0, 0.01, p(d1), p(v1)))
    for i in range(P):
        assert np.array_equal(age1[i], ora.increment_age(age0[i], cam, cam, T10[i], depth0[i]))
        ed, ev = ora.propagate(T10[i], cam, cam, depth0[i], var0[i], 7.0, 50.0, 0.01)
        assert np.array_equal(d1[i], ed) and np.array_equal(v1[i], ev)
    flag = np.where(rng.random((P, H, W)) < 0.5, 0, -6).astype(np.int64)
    mu, var = np.empty((P, H, W)), np.empty((P, H, W))
    _lib.check(_lib.fusion_host(p(depth0), p(d1), p(var0), p(v1), p(flag), depth0.size, 0, p(mu), p(var)))
    emu, evar = ora.fusion_arrays(depth0, d1, var0, v1)
    assert np.array_equal(mu, np.where(flag == 0, emu, depth0)) and np.array_equal(var, np.where(flag == 0, evar, var0))
    reg = np.empty((P, H, W))
    _lib.check(_lib.regularize_host(p(depth0), p(var0), p(flag), P, H, W, p(reg)))
    for i in range(P):
        assert np.array_equal(reg[i], ora.regularize(depth0[i], var0[i], flag[i]))
    assert _lib.regularize_host(None, p(var0), p(flag), P, H, W, p(reg)) == _lib.ERR_INVALID


def test_estimate_debug_outside_the_image_returns_the_reference_flag():
    """src/py/semi_dense.rs:126-155 runs estimate() for any u_key: a pixel outside the image ends at
    KEY_OUT_OF_RANGE unless an earlier test (-7, -1, -8) fails first, and the prior is echoed."""
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import Frame, Params, estimate_debug_
    it = scene.make_pair(77, H=48, W=64, cam=np.array([60., 60., 31.5, 23.5]))
    cp = CameraParameters((60., 60.), (31.5, 23.5))
    key, ref = Frame(cp, it["key_img"], it["T_wk"]), Frame(cp, it["ref_imgs"][0], it["T_wr"][0])
    pr = it["params"]
    params = Params(*pr)
    for u in [(-1, 5), (64, 5), (10, -3), (10, 48), (500, 700)]:
        for d, v in [(200.0, 0.1), (-5.0, 0.1), (1e-9, 1e-3)]:
            want = ora.estimate_debug(u, d, v, it["key_img"], it["key_cam"], it["T_wk"], it["ref_imgs"][0],
                                      it["ref_cams"][0], it["T_wr"][0], ora.make_params(*pr))
            got = estimate_debug_(np.array(u, dtype=np.int64), d, v, key, ref, params)
            assert got == want and got[0] == d and got[1] == v and got[2] in (-2, -7, -1, -8), (u, d, v, got, want)


def test_drop_in_calls_from_several_threads_at_once():
    """Four Python threads call the numpy drop-in at the same time (ctypes releases the GIL): calls
    on one device are serialised inside the library, the worker pool and the pinned rings are
    shared -- every thread gets the oracle's maps for ITS keyframe."""
    import threading
    from rust_bindings.camera import CameraParameters
    from rust_bindings.semi_dense import Frame, Params, increment_age, propagate, update_depth
    from tadataka_b200 import scene
    items = [scene.make_keyframe(640 + i, R=3, H=240, W=320, cam=scene.CAM_VGA / 2) for i in range(4)]
    want, errors, got = [], [], [None] * 4
    for it in items:
        b = scene.stack_keyframes([it])
        want.append(ora.update_depth_batch(b["key_img"], b["key_cam"], b["T_wk"], b["ref_imgs"], b["ref_cams"],
                                           b["T_wr"], b["age"], b["prior_depth"], b["prior_variance"],
                                           ora.make_params(*b["params"]))[:3])

    def work(i):
        try:
            it = items[i]
            cam = CameraParameters((it["key_cam"][0], it["key_cam"][1]), (it["key_cam"][2], it["key_cam"][3]))
            key = Frame(cam, it["key_img"], it["T_wk"])
            refs = [Frame(cam, it["ref_imgs"][r], it["T_wr"][r]) for r in range(3)]
            params = Params(*[float(x) for x in it["params"]])
            T10 = np.linalg.inv(it["T_wk"]) @ it["T_wr"][-1]
            for _ in range(5):
                d, v, f = update_depth(key, refs, it["age"], it["prior_depth"], it["prior_variance"], params)
                a1 = increment_age(it["age"], cam, cam, T10, it["prior_depth"])
                d1, v1 = propagate(T10, cam, cam, it["prior_depth"], it["prior_variance"], 200.0, 100.0, 1e-6)
            got[i] = (d, v, f, a1, d1, v1)
        except Exception as e:   # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errors, errors
    for i, it in enumerate(items):
        d, v, f, a1, d1, v1 = got[i]
        assert np.array_equal(f, want[i][2][0]) and np.array_equal(d, want[i][0][0]) and np.array_equal(v, want[i][1][0])
        cam = it["key_cam"]
        T10 = np.linalg.inv(it["T_wk"]) @ it["T_wr"][-1]
        assert np.array_equal(a1, ora.increment_age(it["age"], cam, cam, T10, it["prior_depth"]))
        od1, ov1 = ora.propagate(T10, cam, cam, it["prior_depth"], it["prior_variance"], 200.0, 100.0, 1e-6)
        assert np.array_equal(d1, od1) and np.array_equal(v1, ov1)
