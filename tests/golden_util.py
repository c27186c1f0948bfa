"""Loader for the committed golden fixtures (tests/golden/)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rgb2gray(rgb):
    """skimage.color.rgb2gray on a uint8 RGB image (what the reference's tests feed the path,
    tests/vo/semi_dense/test_semi_dense.py:55-56), restated because skimage is not installed."""
    f = rgb.astype(np.float64) / 255.
    return np.ascontiguousarray(f @ np.array([0.2125, 0.7154, 0.0721]))


def load_newtsukuba():
    z = np.load(os.path.join(GOLDEN, "newtsukuba_pair0.npz"))
    g = {k: z[k] for k in z.files}
    g["key"] = rgb2gray(g["left_rgb"])
    g["ref"] = rgb2gray(g["right_rgb"])
    return g
