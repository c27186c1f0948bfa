"""tdk_update_depth_static_dev: the update for callers that are not allowed to allocate.

A tracker that must stay clear of cudaMalloc once its loop runs (the call synchronises the whole
device and can fail), or a host program that owns all device memory and gives a plug-in none,
cannot supply the workspace tdk_update_depth_dev asks for.  For them the library carries a small
static device array and this entry point uses it instead.  The test drives it through the C-ABI
exactly as such a caller would -- inputs and outputs are the caller's, nothing else is passed --
and requires bit-identical results to the regular entry point on consecutive calls that share
the array, including a call that follows one which left an error in the status words.

Named test_zz_* so that pytest collects it after every other test file.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from tadataka_b200 import scene

pytestmark = pytest.mark.gpu

NAMES = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth",
         "prior_variance"]


def _batch(seed, P, R, H=96, W=128):
    cam = np.array([100., 100., (W - 1) / 2., (H - 1) / 2.])
    return scene.stack_keyframes([scene.make_keyframe(seed + i, R=R, H=H, W=W, cam=cam)
                                  for i in range(P)])


def _device_inputs(b):
    return {k: torch.from_numpy(np.ascontiguousarray(b[k])).cuda() for k in NAMES}


def _regular(t, params, fuse):
    from tadataka_b200 import device
    d, v, f, am = device.update_depth(*(t[k] for k in NAMES), params, debug=True, fuse=fuse)
    torch.cuda.synchronize()
    return d, v, f, am


def _outputs(t):
    P, H, W = t["key_img"].shape
    return (torch.empty((P, H, W), dtype=torch.float64, device="cuda"),
            torch.empty((P, H, W), dtype=torch.float64, device="cuda"),
            torch.empty((P, H, W), dtype=torch.int64, device="cuda"),
            torch.empty((P, H, W), dtype=torch.int32, device="cuda"))


def _static(t, params, fuse, wait_and_check=1, out=None):
    """Call the new entry point the way a no-allocation caller does: its own input and output
    buffers, the stream, and nothing else."""
    from tadataka_b200 import _lib
    P, H, W = t["key_img"].shape
    R = t["ref_imgs"].shape[1]
    d, v, f, am = out if out is not None else _outputs(t)
    pa = np.ascontiguousarray(params, dtype=np.float64)
    rc = _lib.update_depth_static_dev(
        *(C.c_void_p(t[k].data_ptr()) for k in NAMES), pa.ctypes.data_as(C.c_void_p), P, R, H, W,
        C.c_void_p(d.data_ptr()), C.c_void_p(v.data_ptr()), C.c_void_p(f.data_ptr()),
        C.c_void_p(am.data_ptr()), None, int(fuse), int(wait_and_check),
        C.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return rc, (d, v, f, am)


def _same(a, b):
    return all(torch.equal(x, y) for x, y in zip(a, b))


def test_two_calls_in_a_row_match_the_regular_entry_point():
    from tadataka_b200 import _lib
    first, second = _batch(900, P=2, R=3), _batch(950, P=4, R=8, H=64, W=192)
    t1, t2 = _device_inputs(first), _device_inputs(second)
    want1 = _regular(t1, first["params"], fuse=False)
    want2 = _regular(t2, second["params"], fuse=True)
    rc, got1 = _static(t1, first["params"], fuse=False)
    assert rc == _lib.OK, _lib.last_error()
    rc, got2 = _static(t2, second["params"], fuse=True)      # same array, other shapes and data
    assert rc == _lib.OK, _lib.last_error()
    assert (want1[2] == 0).sum() > 0 and (want2[2] == 0).sum() > 0
    assert _same(got1, want1), "first call differs from tdk_update_depth_dev"
    assert _same(got2, want2), "second call differs from tdk_update_depth_dev"


def test_matches_the_oracle_and_leaves_device_memory_alone():
    """The entry point against the CPU oracle directly (not only against the other CUDA entry
    point), and the property it exists for: once the caller's own buffers are there, calls of
    different shapes change nothing in the amount of free device memory."""
    import oracle as ora
    from tadataka_b200 import _lib
    batches = [_batch(960, P=3, R=4), _batch(970, P=1, R=8, H=80, W=200), _batch(980, P=8, R=8, H=48, W=64)]
    ins = [_device_inputs(b) for b in batches]
    outs = [_outputs(t) for t in ins]
    rc, _ = _static(ins[0], batches[0]["params"], fuse=False, out=outs[0])   # first call: address look-up
    assert rc == _lib.OK, _lib.last_error()
    torch.cuda.synchronize()
    free_before, _total = torch.cuda.mem_get_info()
    for b, t, o in zip(batches, ins, outs):
        rc, _ = _static(t, b["params"], fuse=False, out=o)
        assert rc == _lib.OK, _lib.last_error()
    torch.cuda.synchronize()
    free_after, _total = torch.cuda.mem_get_info()
    assert free_after == free_before, "tdk_update_depth_static_dev changed the free device memory"
    for b, (d, v, f, am) in zip(batches, outs):
        od, ov, of, oam, _ = ora.update_depth_batch(*(b[k] for k in NAMES), ora.make_params(*b["params"]),
                                                    debug=True)
        assert np.array_equal(f.cpu().numpy(), of) and np.array_equal(am.cpu().numpy(), oam)
        assert np.array_equal(d.cpu().numpy(), od, equal_nan=True)
        assert np.array_equal(v.cpu().numpy(), ov, equal_nan=True)
        assert (of == 0).sum() > 0


def test_an_error_left_by_one_call_does_not_leak_into_the_next():
    from tadataka_b200 import _lib
    bad, good = _batch(910, P=1, R=2), _batch(920, P=1, R=2)
    bad["age"][0, 5, 5] = 7                                   # age > R
    rc, _ = _static(_device_inputs(bad), bad["params"], fuse=False)
    assert rc == _lib.ERR_AGE and b"Age exceeds" in _lib.last_error()
    tg = _device_inputs(good)
    rc, got = _static(tg, good["params"], fuse=False)
    assert rc == _lib.OK, _lib.last_error()
    assert _same(got, _regular(tg, good["params"], fuse=False))
    # enqueue-only form: no synchronisation inside, same results once the stream has drained
    rc, got = _static(tg, good["params"], fuse=False, wait_and_check=0)
    assert rc == _lib.OK and _same(got, _regular(tg, good["params"], fuse=False))


def test_batches_beyond_the_built_in_scratch_are_refused():
    from tadataka_b200 import _lib
    b = _batch(930, P=1, R=1, H=32, W=64)
    t = _device_inputs(b)
    fake = dict(t)
    # claim 13 x 5 = 65 pairs: rejected before anything is launched or dereferenced
    P, R, H, W = 13, 5, 32, 64
    pa = np.ascontiguousarray(b["params"], dtype=np.float64)
    z = C.c_void_p(t["key_img"].data_ptr())
    rc = _lib.update_depth_static_dev(z, z, z, z, z, z, z, z, z, pa.ctypes.data_as(C.c_void_p), P, R,
                                      H, W, z, z, z, None, None, 0, 1, None)
    assert rc == _lib.ERR_INVALID
    msg = _lib.last_error().decode()
    assert "at most 64" in msg and "tdk_update_depth_dev" in msg
    assert _lib.STATIC_MAX_PAIRS == 64
    # sizes whose product overflows 64 bits are refused by the per-factor test, not by the product
    rc = _lib.update_depth_static_dev(z, z, z, z, z, z, z, z, z, pa.ctypes.data_as(C.c_void_p), 2 ** 62, 4,
                                      H, W, z, z, z, None, None, 0, 1, None)
    assert rc == _lib.ERR_INVALID
    del fake
