"""Multi-GPU plumbing: keyframes are independent (each carries its own references and maps,
src/semi_dense/semi_dense.rs:160-234), so P keyframes are split into contiguous blocks, one per
rank, and nothing is exchanged until the results are gathered.  torch.distributed only
(NCCL on GPUs, gloo in the CPU tests); never called inside the per-step loop.
"""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous block [start, stop) of rank `rank`; the first n_items % world ranks get one more."""
    if world <= 0 or not 0 <= rank < world or n_items < 0:
        raise ValueError("bad shard arguments")
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def gather_blocks(local, n_items, group=None):
    """All-gather per-rank blocks [n_local, ...] (n_local from shard_range) into [n_items, ...] on
    every rank.  Uneven blocks are padded to the largest one for the collective."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    start, stop = shard_range(n_items, rank, world)
    if local.shape[0] != stop - start:
        raise ValueError(f"rank {rank} holds {local.shape[0]} items, expected {stop - start}")
    biggest = (n_items + world - 1) // world
    if n_items % world == 0:
        out = torch.empty((n_items,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    padded = torch.zeros((biggest,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[:local.shape[0]] = local
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    pieces = []
    for r in range(world):
        a, b = shard_range(n_items, r, world)
        pieces.append(parts[r][:b - a])
    return torch.cat(pieces, dim=0)


def reduce_histogram(hist, group=None):
    """Sum the [.., 10] flag histograms over ranks (in place) and return it."""
    dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    return hist
