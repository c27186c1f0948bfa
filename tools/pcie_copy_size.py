"""Host <-> device copy rate of one GPU against the size of a copy: H2D alone, D2H alone, and both
directions at once, from pinned memory, copies queued back to back on one stream per direction
(and on two streams per direction).  Explains what the chunked host-pointer entry points can reach.
    python tools/pcie_copy_size.py
Development tool."""
import json
import time

import torch

total = 1 << 29
host_a = torch.empty(total, dtype=torch.uint8).pin_memory(); host_a.fill_(1)
host_b = torch.empty(total, dtype=torch.uint8).pin_memory()
dev_a = torch.empty(total, dtype=torch.uint8, device="cuda")
dev_b = torch.ones(total, dtype=torch.uint8, device="cuda")
streams = [torch.cuda.Stream() for _ in range(4)]


def run(size, mode, lanes):
    n = total // size
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        sl = slice(i * size, (i + 1) * size)
        if mode in ("h2d", "both"):
            with torch.cuda.stream(streams[i % lanes]):
                dev_a[sl].copy_(host_a[sl], non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(streams[2 + i % lanes]):
                host_b[sl].copy_(dev_b[sl], non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return (2 if mode == "both" else 1) * n * size / dt / 1e9


out = {}
for size in (307200, 1228800, 2457600, 4915200, 9830400, 39321600, 1 << 28):
    row = {}
    for mode in ("h2d", "d2h", "both"):
        for lanes in (1, 2):
            run(size, mode, lanes)
            row[f"{mode}_{lanes}"] = round(max(run(size, mode, lanes) for _ in range(2)), 1)
    out[size] = row
    print(size, row, flush=True)
print(json.dumps(out))
