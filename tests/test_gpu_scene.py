"""Device ray-caster (csrc/scene.cu) against the numpy renderer it mirrors, and
the device-built keyframe batch through update_depth against the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("pset_name", ["S", "T"])
def test_device_render_matches_numpy_render(pset_name):
    from tadataka_b200 import scene, scene_device
    pset = scene.SET_S if pset_name == "S" else scene.SET_T
    H, W = 240, 320
    cam = scene.CAM_VGA / 2
    world = scene.World(pset, 77)
    b = pset["baseline"]
    poses = np.stack([scene.pose(),
                      scene.pose(scene.rotation(0.004, -0.015, 0.002), (b, -0.1 * b, 0.05 * b)),
                      scene.pose(scene.rotation(-0.003, 0.01, -0.004), (-1.5 * b, 0.12 * b, -0.08 * b))])
    img, dep = scene_device.render(world, poses, np.tile(cam, (3, 1)), H, W)
    img, dep = img.cpu().numpy(), dep.cpu().numpy()
    for i in range(3):
        want_img, want_dep = world.render(poses[i], cam, H, W)
        # sin() and the dot products round differently on the two sides; a pixel whose ray grazes a
        # quad's edge may pick the other plane
        close = (np.abs(img[i] - want_img) <= 1e-9) & (np.abs(dep[i] - want_dep) <= 1e-9 * np.abs(want_dep))
        assert close.mean() >= 0.9999, f"view {i}: {(~close).sum()} pixels differ"
        assert img[i].min() >= 0. and img[i].max() <= 1.
    assert img.std() > 0.01   # textured, not flat


def test_device_built_keyframes_run_through_update_depth_like_the_oracle():
    import torch
    import oracle as ora
    from tadataka_b200 import device, scene, scene_device
    H, W, R = 120, 160, 3
    b = scene_device.make_keyframes([5001, 5002], R=R, H=H, W=W, cam=scene.CAM_VGA / 4)
    assert b["key_img"].shape == (2, H, W) and b["ref_imgs"].shape == (2, R, H, W)
    assert int(b["age"].max()) <= R and float(b["prior_variance"].min()) >= 1e-3
    names = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth", "prior_variance"]
    t = {k: b[k] for k in names}
    d, v, f = device.update_depth(*(t[k] for k in names), b["params"])
    torch.cuda.synchronize()
    host = {k: t[k].cpu().numpy() for k in names}
    host["age"] = host["age"].astype(np.uint64)
    od, ov, of = ora.update_depth_batch(*(host[k] for k in names), ora.make_params(*b["params"]), threads=2)
    assert np.array_equal(f.cpu().numpy(), of)
    assert np.array_equal(d.cpu().numpy(), od, equal_nan=True) and np.array_equal(v.cpu().numpy(), ov, equal_nan=True)
    assert (of == 0).sum() > 50
