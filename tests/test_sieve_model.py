"""The certified sieve (csrc/update_depth.cu, phase A0) decides most pixels from an f32 approximation
with a proven error bound.  tools/sieve_model.py restates that decision in numpy; here the
restatement is checked against the oracle on the CPU: whatever it decides, the oracle must agree
with, and it must decide nearly everything that is not a search candidate."""
import os
import sys

import numpy as np
import pytest

import oracle as ora
from tadataka_b200 import scene

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from sieve_model import cheap_sieve  # noqa: E402


def _cases():
    cam = np.array([260.0, 259.0, 159.5, 119.5])
    yield "C2 blocks", scene.make_keyframe(2100, R=4, H=240, W=320, cam=cam)
    yield "C1 pair", scene.make_pair(1100, H=240, W=320, cam=cam)
    yield "set T", scene.make_pair(4100, H=180, W=320, cam=np.array([225.0, 225.0, 159.5, 89.5]), pset=scene.SET_T)
    it = scene.make_pair(1101, H=120, W=160, cam=np.array([130.0, 130.0, 79.5, 59.5]))
    it["key_img"] = it["key_img"].copy()
    it["key_img"][40:44, 50:60] = np.nan            # a NaN in the image: nothing may be certified from it
    yield "NaN patch", it
    it = scene.make_pair(1102, H=120, W=160, cam=np.array([130.0, 130.0, 79.5, 59.5]))
    it["key_img"] = it["key_img"] * 255.0           # 8-bit intensity scale with the same threshold
    it["ref_imgs"] = it["ref_imgs"] * 255.0
    yield "0..255 scale", it


@pytest.mark.parametrize("name,it", list(_cases()), ids=[n for n, _ in _cases()])
def test_certified_decisions_never_contradict_the_oracle(name, it):
    b = scene.stack_keyframes([it])
    _, _, f = ora.update_depth_batch(b["key_img"], b["key_cam"], b["T_wk"], b["ref_imgs"], b["ref_cams"], b["T_wr"],
                                     b["age"], b["prior_depth"], b["prior_variance"], ora.make_params(*b["params"]))[:3]
    f = f[0]
    m = cheap_sieve(it["key_img"], it["key_cam"], it["T_wk"], it["ref_cams"], it["T_wr"], it["age"],
                    it["prior_depth"], it["prior_variance"], it["params"])
    decided = m != 0
    assert not np.any(decided & (m != f)), f"{name}: a certified decision differs from the oracle"
    if name in ("C2 blocks", "C1 pair", "set T"):
        # undecided = the search candidates plus a thin band near the threshold / the border
        not_candidates = (f == -6) | (f == -2)
        left = (~decided & not_candidates).sum() / max(not_candidates.sum(), 1)
        assert left < 0.05, f"{name}: {left:.3f} of the non-candidates stayed undecided"
