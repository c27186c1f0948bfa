"""End-to-end pin of the oracle's `estimate` on SUCCESS pixels.

The golden maps of tests/golden/newtsukuba_pair0.npz were produced by the oracle itself, so
comparing the GPU with them is GPU-vs-oracle.  What IS pinned by the reference are the primitives
(tests/test_oracle_golden.py replays the reference's own #[test] vectors for each of them).  This
test composes those pinned primitives in Python in the order of src/semi_dense/semi_dense.rs:91-158
-- an independent second writing of `estimate` -- and requires the C oracle's full-pixel result to
equal the composition bit for bit on successful pixels of the reference's NewTsukuba fixture, so the
success path (triangulation, alpha, geo/photo variance, final check) is tied to reference-held
vectors through the primitives rather than only to itself.
"""
import numpy as np

import oracle as ora
from tests.golden_util import load_newtsukuba


def _estimate_by_composition(u, prior_depth, prior_var, g, params6):
    vmin, vmax = ora.safe_invert(params6[1]), ora.safe_invert(params6[0])       # py/semi_dense.rs:102-107
    cam, T_wk, T_wr = g["cam"], g["T_wl"], g["T_wr"]
    T_rk = ora.transform_rk(T_wk, T_wr)                                          # :101
    idp = ora.safe_invert(prior_depth)
    assert ora.check_args(idp, prior_var, (vmin, vmax)) == 0
    lo, hi = ora.hypothesis_range(idp, prior_var, (vmin, vmax))
    depth_range = ora.depth_search_range(lo, hi)                                 # :103
    x_key = ora.cam_normalize(cam, np.array(u, dtype=np.float64))                # :104
    flag, ratio = ora.step_ratio(T_rk, x_key, idp)                               # :108
    assert flag == 0
    key_step = ratio * params6[4]
    x_min, x_max = ora.calc_ref_ends(T_rk, x_key, depth_range)                   # :115
    ref_dir = x_max - x_min
    e_key = ora.calc_key_epipole(T_wk, T_wr)                                     # :118
    key_dir = ora.calc_key_direction(x_key, e_key, ref_dir)
    xs_key = ora.key_coordinates(key_dir, x_key, key_step)                       # :122
    us_key = np.stack([ora.cam_unnormalize(cam, x) for x in xs_key])
    H, W = g["key"].shape
    assert all(ora.is_in_range(x, y, (H, W)) for x, y in us_key)
    ki = ora.interpolate(g["key"], us_key)                                       # :129
    key_gradient = ora.norm(ora.gradient1d(ki))
    assert not key_gradient < params6[5]
    xs_ref = ora.ref_coordinates(x_min, ref_dir, params6[4])                     # :139
    us_ref = np.stack([ora.cam_unnormalize(cam, x) for x in xs_ref])
    assert ora.check_us_ref(us_ref, 5, (H, W)) == 0
    ri = ora.interpolate(g["ref"], us_ref)                                       # :144
    argmin = ora.search(ri, ki)                                                  # :147
    key_depth = ora.calc_depth0(T_rk, x_key, xs_ref[argmin])                     # :148
    alpha = ora.calc_alpha(T_rk, x_key, depth_range, key_depth)                  # :150
    gx, gy = ora.sobel(g["key"])
    grad = np.array([gx[u[1], u[0]], gy[u[1], u[0]]])
    geo = ora.geo_var(x_key, T_rk[:3, 3], grad)                                  # :152
    photo = ora.photo_var(key_gradient / key_step)                               # :153
    variance = ora.calc_variance(alpha, geo, photo, params6[2], params6[3])
    idn = ora.safe_invert(key_depth)
    assert ora.check_args(idn, variance, (vmin, vmax)) == 0                      # :156
    return ora.safe_invert(idn), variance, int(argmin)                           # update_depth stores inv(inv_depth)


def test_success_pixels_equal_the_composition_of_pinned_primitives():
    g = load_newtsukuba()
    params6 = (60.0, 1000.0, 0.01, 0.01, 0.01, 0.2)
    p = ora.make_params(*params6)
    ys, xs = np.nonzero(g["flag_map"] == 0)
    assert ys.size > 1000
    picks = np.linspace(0, ys.size - 1, 25).astype(int)
    for k in picks:
        u = (int(xs[k]), int(ys[k]))
        d, v, f = ora.estimate_debug(u, 200.0, 1.0, g["key"], g["cam"], g["T_wl"], g["ref"], g["cam"], g["T_wr"], p)
        assert f == 0
        cd, cv, _ = _estimate_by_composition(u, 200.0, 1.0, g, params6)
        assert d == cd and v == cv, (u, d, cd, v, cv)
