"""Two (or more) GPUs: keyframes sharded over the ranks, results gathered on rank 0 over NCCL
(tadataka_b200.sharding.RootGather) -- the gathered maps must equal a single-GPU run bit for bit.
Skipped on a box with one GPU (the gloo version of the same logic runs in test_sharding.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

H, W, N_PAIRS, CHUNK = 240, 320, 18, 4
CAM = np.array([260.0, 259.0, 159.5, 119.5])
NAMES = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth", "prior_variance"]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    from tadataka_b200 import device, scene_device, sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        rg = sharding.RootGather(N_PAIRS, H, W, CHUNK, dev)
        hist = torch.zeros((CHUNK, 10), dtype=torch.int64, device=dev)
        params = None
        for j in range(rg.rounds):
            a, z = rg.local_span(j)
            views = rg.out_views(j)
            if z > a:
                t = scene_device.make_keyframes([5000 + i for i in range(rg.start + a, rg.start + z)], R=2, H=H, W=W,
                                                cam=CAM, device=dev, per_item_rng=True)
                params = t["params"]
                device.update_depth(*(t[k] for k in NAMES), params, check=True, out=views, hist=hist[:z - a])
            rg.ship(j)
        rg.finish()
        total = sharding.reduce_histogram(hist.sum(dim=0))
        torch.cuda.synchronize()
        if rank == 0:
            # the same pairs in one launch on one GPU
            t = scene_device.make_keyframes([5000 + i for i in range(N_PAIRS)], R=2, H=H, W=W, cam=CAM, device=dev,
                                            per_item_rng=True)
            d, v, f, h = device.update_depth(*(t[k] for k in NAMES), t["params"], hist=True)
            torch.cuda.synchronize()
            np.savez(tmp, depth_ok=bool(torch.equal(d, rg.depth)), var_ok=bool(torch.equal(v, rg.variance)),
                     flag_ok=bool(torch.equal(f, rg.flag)), hist=total.cpu().numpy(), hist_single=h.sum(dim=0).cpu().numpy(),
                     success=int((f == 0).sum()), received=rg.bytes_received)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_sharded_run_gathered_on_rank0_equals_single_gpu(tmp_path):
    world = min(torch.cuda.device_count(), 4)
    out = str(tmp_path / "multi.npz")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    assert got["depth_ok"] and got["var_ok"] and got["flag_ok"]
    assert got["hist"].tolist() == got["hist_single"].tolist()
    assert int(got["hist"].sum()) == N_PAIRS * H * W and int(got["success"]) > 0
    assert int(got["received"]) > 0
