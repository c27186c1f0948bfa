"""Host <-> device copy rates of every GPU of the box, one rank per GPU, alone and all at once.
What limits the end-to-end (host-buffer) numbers at N > 1: the PCIe link of a GPU, a switch shared by
several GPUs, or the host memory the pinned buffers live in.
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py
Prints one JSON line (rank 0).  Development tool."""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
info = {"rank": rank}
try:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(local)
    pynvml.nvmlDeviceSetCpuAffinity(h)                      # run (and first-touch pinned memory) next to the GPU
    bus = pynvml.nvmlDeviceGetPciInfo(h).busId
    bus = bus.decode() if isinstance(bus, bytes) else bus
    info["bus"] = bus
    try:
        info["numa_node"] = int(open(f"/sys/bus/pci/devices/{bus.lower()[4:] if len(bus) > 12 else bus.lower()}/numa_node").read())
    except Exception:
        info["numa_node"] = None
    info["pcie_gen"] = pynvml.nvmlDeviceGetCurrPcieLinkGeneration(h)
    info["pcie_width"] = pynvml.nvmlDeviceGetCurrPcieLinkWidth(h)
    info["cpus"] = len(os.sched_getaffinity(0))
except Exception as e:                                       # noqa: BLE001
    info["nvml_error"] = str(e)
if world > 1:
    dist.init_process_group("gloo")
n = 1 << 30
host = torch.empty(n, dtype=torch.uint8).pin_memory()
host.fill_(1)
devb = torch.empty(n, dtype=torch.uint8, device="cuda")
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
host2 = torch.empty(n, dtype=torch.uint8).pin_memory()
dev2 = torch.ones(n, dtype=torch.uint8, device="cuda")


def rate(mode, reps=4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s_in):
                devb.copy_(host, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s_out):
                host2.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return (2 if mode == "both" else 1) * n * reps / dt / 1e9


def barrier():
    if world > 1:
        dist.barrier()


res = {}
for mode in ("h2d", "d2h", "both"):
    rate(mode, 1)
    # alone: ranks take turns
    alone = None
    for r in range(world):
        barrier()
        if r == rank:
            alone = rate(mode)
        barrier()
    barrier()
    together = rate(mode)
    barrier()
    res[mode] = {"alone_gbs": alone, "all_ranks_at_once_gbs": together}
info["rates"] = res
if world > 1:
    out = [None] * world
    dist.all_gather_object(out, info)
else:
    out = [info]
if rank == 0:
    agg = {m: sum(o["rates"][m]["all_ranks_at_once_gbs"] for o in out) for m in ("h2d", "d2h", "both")}
    print(json.dumps({"ranks": out, "aggregate_all_at_once_gbs": agg}))
if world > 1:
    dist.destroy_process_group()
