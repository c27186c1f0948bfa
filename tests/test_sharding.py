"""N > 1 host logic on CPU: world_size-2 gloo, the oracle standing in for the kernel."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tadataka_b200.sharding import shard_range


def test_shard_range_covers_everything_once():
    for n in (0, 1, 7, 64, 4096, 4099):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, tmp):
    import oracle as ora
    from tadataka_b200 import scene
    from tadataka_b200.sharding import gather_blocks, reduce_histogram
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cam = np.array([60., 60., 31.5, 23.5])
        items = [scene.make_keyframe(8000 + i, R=2, H=48, W=64, cam=cam) for i in range(n_items)]
        b = scene.stack_keyframes(items)
        a, z = shard_range(n_items, rank, world)
        sl = slice(a, z)
        d, v, f = ora.update_depth_batch(b["key_img"][sl], b["key_cam"][sl], b["T_wk"][sl],
                                         b["ref_imgs"][sl], b["ref_cams"][sl], b["T_wr"][sl],
                                         b["age"][sl], b["prior_depth"][sl], b["prior_variance"][sl],
                                         ora.make_params(*b["params"]))
        depth_all = gather_blocks(torch.from_numpy(d), n_items)
        flag_all = gather_blocks(torch.from_numpy(f), n_items)
        hist = torch.tensor([int((f == -k).sum()) for k in range(10)], dtype=torch.int64)
        reduce_histogram(hist)
        if rank == 0:
            np.savez(tmp, depth=depth_all.numpy(), flag=flag_all.numpy(), hist=hist.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_items", [4, 5])
def test_two_ranks_gather_equals_single_process(tmp_path, n_items):
    import oracle as ora
    from tadataka_b200 import scene
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(2, _free_port(), n_items, out), nprocs=2, join=True)
    got = np.load(out)
    cam = np.array([60., 60., 31.5, 23.5])
    b = scene.stack_keyframes([scene.make_keyframe(8000 + i, R=2, H=48, W=64, cam=cam)
                               for i in range(n_items)])
    d, v, f = ora.update_depth_batch(b["key_img"], b["key_cam"], b["T_wk"], b["ref_imgs"],
                                     b["ref_cams"], b["T_wr"], b["age"], b["prior_depth"],
                                     b["prior_variance"], ora.make_params(*b["params"]))
    assert np.array_equal(got["depth"], d) and np.array_equal(got["flag"], f)
    assert got["hist"].tolist() == [int((f == -k).sum()) for k in range(10)]


def _root_gather_worker(rank, world, port, n_items, chunk, tmp):
    import oracle as ora
    from tadataka_b200 import scene
    from tadataka_b200.sharding import RootGather, reduce_histogram
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cam = np.array([60., 60., 31.5, 23.5])
        b = scene.stack_keyframes([scene.make_keyframe(8100 + i, R=2, H=48, W=64, cam=cam)
                                   for i in range(n_items)])
        rg = RootGather(n_items, 48, 64, chunk, "cpu")
        hist = torch.zeros(10, dtype=torch.int64)
        for j in range(rg.rounds):
            a, z = rg.local_span(j)
            od, ov, of = rg.out_views(j)
            if z > a:
                sl = slice(rg.start + a, rg.start + z)
                d, v, f = ora.update_depth_batch(b["key_img"][sl], b["key_cam"][sl], b["T_wk"][sl],
                                                 b["ref_imgs"][sl], b["ref_cams"][sl], b["T_wr"][sl],
                                                 b["age"][sl], b["prior_depth"][sl], b["prior_variance"][sl],
                                                 ora.make_params(*b["params"]))
                od.copy_(torch.from_numpy(d)); ov.copy_(torch.from_numpy(v)); of.copy_(torch.from_numpy(f))
                hist += torch.tensor([int((f == -k).sum()) for k in range(10)])
            rg.ship(j)
        rg.finish()
        reduce_histogram(hist)
        if rank == 0:
            np.savez(tmp, depth=rg.depth.numpy(), variance=rg.variance.numpy(), flag=rg.flag.numpy(),
                     hist=hist.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_items,chunk", [(5, 1), (6, 2), (3, 4)])
def test_root_gather_in_rounds_equals_single_process(tmp_path, n_items, chunk):
    """RootGather: two ranks, rounds of `chunk` keyframes, transfers overlapping the next round;
    the root ends up with exactly the maps of a single-process run (flags went over the wire as int8)."""
    import oracle as ora
    from tadataka_b200 import scene
    out = str(tmp_path / "root.npz")
    mp.spawn(_root_gather_worker, args=(2, _free_port(), n_items, chunk, out), nprocs=2, join=True)
    got = np.load(out)
    cam = np.array([60., 60., 31.5, 23.5])
    b = scene.stack_keyframes([scene.make_keyframe(8100 + i, R=2, H=48, W=64, cam=cam)
                               for i in range(n_items)])
    d, v, f = ora.update_depth_batch(b["key_img"], b["key_cam"], b["T_wk"], b["ref_imgs"],
                                     b["ref_cams"], b["T_wr"], b["age"], b["prior_depth"],
                                     b["prior_variance"], ora.make_params(*b["params"]))
    assert np.array_equal(got["depth"], d) and np.array_equal(got["variance"], v)
    assert np.array_equal(got["flag"], f)
    assert got["hist"].tolist() == [int((f == -k).sum()) for k in range(10)]
