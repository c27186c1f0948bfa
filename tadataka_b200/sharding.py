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


class RootGather:
    """Gathers the three result maps of a sharded update_depth run on ONE rank while the other
    ranks keep computing.

    n_items keyframes are split by shard_range; every rank works through its block in rounds of at
    most `chunk` keyframes.  After round j a rank hands that round's depth / variance / flag maps to
    `ship(j)`: the root receives them straight into its [n_items, H, W] result arrays (its own
    rounds are computed in place there), the other ranks send from double-buffered chunk buffers,
    so the transfer of round j overlaps the kernels of round j + 1.  Flags travel as int8 (the ten
    ResultFlag values fit) and are widened to int64 on arrival; depth and variance travel as the
    f64 they are -- the gathered maps equal a single-process run bit for bit.

    One grouped NCCL send/recv per round (torch.distributed.batch_isend_irecv) on a side stream;
    gloo on CPU tensors in the tests.  Nothing here runs inside a kernel's inner loop.
    """

    def __init__(self, n_items, H, W, chunk, device, root=0, group=None, cyclic=False,
                 maps=("depth", "variance", "flag")):
        """cyclic=True: round numbers may run past `rounds` and wrap around (a stream of steps over
        the same n_items keyframes, each step's results replacing the previous step's on the root).
        maps: which of the three result maps travel (the others stay on the rank that made them)."""
        self.group, self.root, self.cyclic = group, root, cyclic
        self.maps = tuple(maps)
        self.bytes_per_px = 8 * ("depth" in self.maps) + 8 * ("variance" in self.maps) + 1 * ("flag" in self.maps)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n_items, self.H, self.W, self.chunk = n_items, H, W, max(1, int(chunk))
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.spans = [shard_range(n_items, r, self.world) for r in range(self.world)]
        self.start, self.stop = self.spans[self.rank]
        self.n_local = self.stop - self.start
        biggest = max(b - a for a, b in self.spans)
        self.rounds = (biggest + self.chunk - 1) // self.chunk
        self.is_root = self.rank == root
        f64, dev = torch.float64, self.device
        if self.is_root:
            self.depth = torch.empty((n_items, H, W), dtype=f64, device=dev)
            self.variance = torch.empty((n_items, H, W), dtype=f64, device=dev)
            self.flag = torch.empty((n_items, H, W), dtype=torch.int64, device=dev)
            # int8 flags of the peers' current round, two sets (one being widened while the next arrives)
            self._flag8 = [torch.empty((self.world, self.chunk, H, W), dtype=torch.int8, device=dev)
                           for _ in range(2)]
        else:
            n = min(self.chunk, max(self.n_local, 1))
            self._buf = [(torch.empty((n, H, W), dtype=f64, device=dev), torch.empty((n, H, W), dtype=f64, device=dev),
                          torch.empty((n, H, W), dtype=torch.int64, device=dev),
                          torch.empty((n, H, W), dtype=torch.int8, device=dev)) for _ in range(2)]
        self._works = [None, None]          # outstanding transfers that use buffer set 0 / 1
        self._pending = []                   # (round, works) on the root: flags still to be widened
        self._comm = torch.cuda.Stream(device=dev) if self.cuda else None
        self.bytes_received = 0

    # ---- what a rank computes in round j
    def local_span(self, j, rank=None):
        """[a, z) within the rank's own block (empty when the block has fewer rounds)."""
        n = self.n_local if rank is None else self.spans[rank][1] - self.spans[rank][0]
        if self.cyclic:
            j = j % self.rounds
        a = min(j * self.chunk, n)
        return a, min(a + self.chunk, n)

    def out_views(self, j):
        """(depth, variance, flag) tensors [z - a, H, W] the round's kernels write into."""
        a, z = self.local_span(j)
        if self.is_root:
            g = slice(self.start + a, self.start + z)
            return self.depth[g], self.variance[g], self.flag[g]
        self._wait(j % 2)                    # the transfer of round j - 2 read this buffer set
        d, v, f, _ = self._buf[j % 2]
        return d[:z - a], v[:z - a], f[:z - a]

    def _wait(self, k):
        if self._works[k] is not None:
            for w in self._works[k]:
                w.wait()
            self._works[k] = None

    # ---- hand round j to the wire (call after the round's kernels were enqueued)
    def ship(self, j):
        if self.world == 1:
            return
        ops = []
        if self.cuda:
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(self.device))
            self._comm.wait_event(ready)
        ctx = torch.cuda.stream(self._comm) if self.cuda else _NullContext()
        with ctx:
            if self.is_root:
                self._widen_done(keep=1)
                self._wait(j % 2)
                for r in range(self.world):
                    if r == self.root:
                        continue
                    a, z = self.local_span(j, r)
                    if z == a:
                        continue
                    g = slice(self.spans[r][0] + a, self.spans[r][0] + z)
                    f8 = self._flag8[j % 2][r, :z - a]
                    if "depth" in self.maps:
                        ops.append(dist.P2POp(dist.irecv, self.depth[g], r, self.group))
                    if "variance" in self.maps:
                        ops.append(dist.P2POp(dist.irecv, self.variance[g], r, self.group))
                    if "flag" in self.maps:
                        ops.append(dist.P2POp(dist.irecv, f8, r, self.group))
                    self.bytes_received += (z - a) * self.H * self.W * self.bytes_per_px
                if ops:
                    works = dist.batch_isend_irecv(ops)
                    self._works[j % 2] = works
                    self._pending.append((j, works))
            else:
                a, z = self.local_span(j)
                if z > a:
                    d, v, f, f8 = self._buf[j % 2]
                    if "depth" in self.maps:
                        ops.append(dist.P2POp(dist.isend, d[:z - a], self.root, self.group))
                    if "variance" in self.maps:
                        ops.append(dist.P2POp(dist.isend, v[:z - a], self.root, self.group))
                    if "flag" in self.maps:
                        f8[:z - a].copy_(f[:z - a])      # int64 -> int8 on the side stream
                        ops.append(dist.P2POp(dist.isend, f8[:z - a], self.root, self.group))
                    if ops:
                        self._works[j % 2] = dist.batch_isend_irecv(ops)

    def _widen_done(self, keep=0):
        """Root: widen the int8 flags of every round whose transfer was issued at least `keep`
        rounds ago (waits for that transfer on the side stream)."""
        while len(self._pending) > keep:
            j, works = self._pending.pop(0)
            for w in works:
                w.wait()
            for r in range(self.world):
                if r == self.root:
                    continue
                a, z = self.local_span(j, r)
                if z > a and "flag" in self.maps:
                    g = slice(self.spans[r][0] + a, self.spans[r][0] + z)
                    self.flag[g].copy_(self._flag8[j % 2][r, :z - a])
            self._works[j % 2] = None

    def finish(self):
        """Wait for every transfer (and the last widening); afterwards the root's depth / variance /
        flag arrays hold all n_items keyframes.  Makes the caller's current stream wait for the side
        stream; does not synchronise the host."""
        if self.world == 1:
            return
        ctx = torch.cuda.stream(self._comm) if self.cuda else _NullContext()
        with ctx:
            if self.is_root:
                self._widen_done(keep=0)
            else:
                self._wait(0)
                self._wait(1)
        if self.cuda:
            done = torch.cuda.Event()
            done.record(self._comm)
            torch.cuda.current_stream(self.device).wait_event(done)


class _NullContext:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False
