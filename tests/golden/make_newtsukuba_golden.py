"""Generates tests/golden/newtsukuba_pair0.npz from the reference's own test fixture.

Run in the build container only (needs /root/reference):
    python tests/golden/make_newtsukuba_golden.py

Stores the frame-0 stereo pair of tests/dataset/new_tsukuba (RGB uint8, alpha dropped) with the
poses and intrinsics `NewTsukubaDataset` would produce (tadataka/dataset/new_tsukuba.py:56-101,
153-158), so that the replayable cases of tests/vo/semi_dense/test_semi_dense.py:104-134 and the
`test_update_depth` smoke (:40-73) can be re-run on the GPU box, where /root/reference is absent.
The oracle's outputs on that pair are stored next to the inputs as the golden answers.

Note: the maps stored here come from the repo's own oracle, not from a build of the reference (no
Rust toolchain): they pin GPU-vs-oracle.  The reference-held evidence is the primitives' vectors
(tests/test_oracle_golden.py), the five flag cases and the flag histogram; tests/test_oracle_chain.py
ties the oracle's success path to those primitives.
"""
import os
import sys

import cv2
import numpy as np
from scipy.spatial.transform import Rotation

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle as ora  # noqa: E402
from tests.golden_util import rgb2gray  # noqa: E402

ROOT = "/root/reference/tests/dataset/new_tsukuba"


def load_rgb(path):
    bgra = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    return np.ascontiguousarray(bgra[:, :, 2::-1])  # -> RGB, alpha dropped (new_tsukuba.py:85-86)


def poses():
    track = np.loadtxt(os.path.join(ROOT, "groundtruth", "camera_track.txt"), delimiter=",")
    positions, euler = track[:, 0:3].copy(), track[:, 3:6].copy()
    R180 = Rotation.from_rotvec([np.pi, 0, 0]).as_matrix()
    positions = (R180 @ positions.T).T
    euler[:, 1:3] = -euler[:, 1:3]
    rot = Rotation.from_euler("xyz", euler, degrees=True)
    return rot, positions


def motion_matrix(R, t):
    T = np.eye(4)
    T[:3, :3], T[:3, 3] = R, t
    return T


def main():
    rot, pos = poses()
    R0 = rot[0].as_matrix()
    offset = R0 @ np.array([10.0, 0, 0])
    T_wl = motion_matrix(R0, pos[0] - offset / 2.0)
    T_wr = motion_matrix(R0, pos[0] + offset / 2.0)
    left = load_rgb(os.path.join(ROOT, "illumination/daylight/left/tsukuba_daylight_L_00201.png"))
    right = load_rgb(os.path.join(ROOT, "illumination/daylight/right/tsukuba_daylight_R_00201.png"))
    cam = ora.cam((615., 615.), (320., 240.))
    key, ref = rgb2gray(left), rgb2gray(right)

    # tests/vo/semi_dense/test_semi_dense.py:104-134 -- the five replayable cases
    p_est = ora.make_params(0.1, 1000.0, 0.01, 0.01, 0.01, 0.2)
    cases = [((110, 400), -10.0, 10.0, -7), ((110, 400), 0.05, 0.2, -1), ((390, 100), 2.0, 0.2, -6),
             ((0, 200), 2.0, 0.2, -2), ((116, 400), 2.0, 0.001, -5)]
    est = []
    for u, d, v, want in cases:
        dd, vv, flag = ora.estimate_debug(u, d, v, key, cam, T_wl, ref, cam, T_wr, p_est)
        assert flag == want, (u, flag, want)
        est.append([u[0], u[1], d, v, flag, dd, vv])

    # tests/vo/semi_dense/test_semi_dense.py:40-73 -- full-map smoke, now with values
    p_upd = ora.make_params(60.0, 1000.0, 0.01, 0.01, 0.01, 0.2)
    H, W = key.shape
    age = np.ones((H, W), dtype=np.uint64)
    d, v, f, am, nn = ora.update_depth(key, cam, T_wl, ref[None], cam[None], T_wr[None], age,
                                       np.full((H, W), 200.0), np.ones((H, W)), p_upd, debug=True)
    hist = {int(k): int((f == k).sum()) for k in range(-9, 1)}
    print("update_depth flag histogram:", hist)
    # an independent restatement got exactly these counts
    assert hist[-6] == 254554 and hist[-2] == 12480 and hist[-3] == 7571 and hist[0] == 32595
    ok = f == 0
    np.savez_compressed(
        os.path.join(HERE, "newtsukuba_pair0.npz"),
        left_rgb=left, right_rgb=right, T_wl=T_wl, T_wr=T_wr, cam=cam,
        estimate_cases=np.array(est), flag_hist=np.array([hist[k] for k in range(-9, 1)]),
        flag_map=f.astype(np.int8), argmin_map=am.astype(np.int16),
        depth_success=d[ok], variance_success=v[ok])
    print("wrote newtsukuba_pair0.npz",
          os.path.getsize(os.path.join(HERE, "newtsukuba_pair0.npz")) / 1e6, "MB")


if __name__ == "__main__":
    main()
