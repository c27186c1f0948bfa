#!/usr/bin/env python
"""bench.py -- semi-dense depth updates (pixels/sec) at 640x480, the metric of BASELINE.json.

Workload (BASELINE.json configs[1], SURVEY 8d "C2"), per GPU and per step:
    B keyframes 640x480, each matched against 8 reference frames (one per pixel, chosen by the
    pixel's age in 32x32 blocks, age 0 = not processed), successful measurements fused with the
    prior by Gaussian fusion in inverse depth.  Synthetic piecewise-planar scenes
    (tadataka_b200/scene.py), f64 images and maps as the reference stores them.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    (N > 1: launched by `python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N`;
     a plain `python bench.py --gpus N` re-launches itself that way.)

Prints ONE JSON line (rank 0).  `value` = device-resident throughput of the C2 step (weak scaling:
every rank runs its own B keyframes per step); at N > 1 every step's depth / variance / flag maps
are gathered on rank 0 over NCCL inside the timed region (tadataka_b200.sharding.RootGather: the
transfer of step k overlaps the kernels of step k + 1; flags travel as int8) and the flag
histograms are all-reduced.  `config.c5` = BASELINE.json configs[4]: 4096 640x480 pairs sharded
over the N ranks (strong scaling), kernels + the same gather, one total time.  `e2e` = the C2 step
through the host-pointer C-ABI call (tdk_update_depth_host) from pinned host buffers, H2D and D2H
inside the timed region; `roofline` = algorithmic bytes of the step / its CUDA-event time vs the
measured HBM peak; `cpu_baseline` = the oracle (CPU restatement of the reference) on the host.
"""
import argparse
import ctypes as C
import json
import os
import socket
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W, R = 480, 640, 8
METRIC = "semi-dense depth updates (pixels/sec) at 640x480"
UNIT = "pixels/s"
# key 8 + one reference 8 + prior depth 8 + prior variance 8 + age 8 read,
# depth 8 + variance 8 + flag 8 written (each pixel uses ONE reference, chosen by age)
ALGO_BYTES_PER_PX = 64


def workload_name(B):
    return (f"C2: {B} x (640x480 keyframe vs 8 reference frames by age + Gaussian fusion) per GPU "
            "per step")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--keyframes", type=int, default=64, help="keyframes per GPU per step")
    ap.add_argument("--distinct", type=int, default=8, help="distinct rendered scenes per GPU")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = min(steps, 30)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c4", action="store_true", help="skip the secondary 1280x720 and keyframe-loop measurements")
    ap.add_argument("--c4-pairs", type=int, default=64, help="1280x720 pairs per launch (BASELINE.json configs[3])")
    ap.add_argument("--c5-pairs", type=int, default=4096, help="640x480 pairs of the sharded run (configs[4]); 0 = skip")
    ap.add_argument("--c5-chunk", type=int, default=64, help="pairs per round of the sharded run")
    ap.add_argument("--nccl-default-priority", action="store_true", help="NCCL on a normal-priority stream (experiment)")
    ap.add_argument("--device-scenes", action="store_true",
                    help="render every keyframe's world on the GPU (tdk_render_scene_dev) instead of in numpy")
    ap.add_argument("--plan", type=int, default=-1, help="kernel plan override (2 sorted [default], 0 two kernels, 1 one kernel)")
    return ap.parse_args()


# --------------------------------------------------------------------------------- workload
def make_workload(rank, B, distinct):
    """B keyframes: `distinct` rendered scenes, each reused for B/distinct keyframes with its own
    seeded age map and priors (every keyframe is a separate copy in memory)."""
    from tadataka_b200 import scene
    distinct = max(1, min(distinct, B))
    base = [scene.make_keyframe(2000 + 1000 * rank + i, R=R, H=H, W=W) for i in range(distinct)]
    items = []
    for k in range(B):
        src = base[k % distinct]
        if k < distinct:
            items.append(src)
            continue
        rng = np.random.default_rng(7_000_000 + 10_000 * rank + k)
        it = dict(src)
        blocks = rng.integers(0, R + 1, size=((H + 31) // 32, (W + 31) // 32))
        it["age"] = np.ascontiguousarray(
            np.kron(blocks, np.ones((32, 32), dtype=np.int64))[:H, :W].astype(np.uint64))
        lo, hi = src["params"][0], src["params"][1]
        it["prior_depth"] = np.clip(src["gt_depth"] * (1.0 + rng.uniform(-0.1, 0.1, size=(H, W))),
                                    lo * 1.01, hi * 0.99)
        it["prior_variance"] = 10.0 ** rng.uniform(-3.0, 0.0, size=(H, W))
        items.append(it)
    return scene.stack_keyframes(items)


def cpu_pipeline(ora, b, sl, threads):
    """The oracle's version of one step on keyframes b[sl]: update_depth + Gaussian fusion."""
    d, v, f = ora.update_depth_batch(b["key_img"][sl], b["key_cam"][sl], b["T_wk"][sl],
                                     b["ref_imgs"][sl], b["ref_cams"][sl], b["T_wr"][sl],
                                     b["age"][sl], b["prior_depth"][sl], b["prior_variance"][sl],
                                     ora.make_params(*b["params"]), threads=threads)
    eps = np.finfo(np.float64).eps
    pd, pv = b["prior_depth"][sl], b["prior_variance"][sl]
    ok = f == 0
    with np.errstate(all="ignore"):
        ia, ib = 1. / (pd[ok] + eps), 1. / (d[ok] + eps)
        vs = pv[ok] + v[ok]
        d[ok] = 1. / ((ia * v[ok] + ib * pv[ok]) / vs + eps)
        v[ok] = (pv[ok] * v[ok]) / vs
    return d, v, f


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def time_cpu(ora, b, n_keyframes, threads, budget_s=20.0, max_reps=5, keep=None):
    """Best of up to max_reps runs of the oracle pipeline on the first n_keyframes keyframes.
    keep (a list) receives the (depth, variance, flag) maps of the last run."""
    sl = slice(0, n_keyframes)
    cpu_pipeline(ora, b, slice(0, min(n_keyframes, threads)), threads)      # warm-up
    best, spent, reps, res = None, 0.0, 0, None
    while reps < max_reps and (reps < 2 or spent < budget_s):
        t0 = time.perf_counter()
        res = cpu_pipeline(ora, b, sl, threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
        spent += dt
        reps += 1
    if keep is not None:
        keep.extend(res)
    return n_keyframes * H * W / best, best, reps


# --------------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""
    BAD = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
           "hw_power_brake_slowdown": 0x80}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, index, period=0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.stop_flag = [], set(), False
        self.max_mhz, self.ok = None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        while not self.stop_flag:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for name, bit in {**self.BAD, **self.NOTE}.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def result(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def profiled_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of update_depth_kernel from the
    committed ncu --set full capture of this same workload (profiles/traffic.json), else None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return t.get("update_depth_kernel_bytes_per_px")
    except Exception:
        return None


# --------------------------------------------------------------------------------- arms
def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path.  The Rust crate cannot be built in this image
    (no cargo/rustc), so this is the oracle port (oracle/tdk_oracle.c), all host threads, on the
    same keyframes as the GPU arm's rank 0 (all B of them per step unless that would run for more
    than about three minutes, in which case every step takes the first S and says so)."""
    if rank != 0:
        return
    import oracle as ora
    threads = host_threads()
    B = args.keyframes
    b = make_workload(0, B, args.distinct)
    t0 = time.perf_counter()
    cpu_pipeline(ora, b, slice(0, min(B, threads)), threads)
    t_unit = (time.perf_counter() - t0) / min(B, threads) * threads   # thread-seconds per keyframe
    total_steps = args.steps + args.warmup
    S = int(max(1, min(B, (180.0 * threads / max(t_unit, 1e-6)) / max(total_steps, 1))))
    if threads <= S < B:
        S = S // threads * threads
    sl = slice(0, S)
    for _ in range(args.warmup):
        cpu_pipeline(ora, b, sl, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pipeline(ora, b, sl, threads)
    dt = time.perf_counter() - t0
    value = S * H * W * args.steps / dt
    sample = (f"{S} of the step's {B} keyframes (640x480, 8 refs by age, fusion), {threads} threads, "
              "OpenMP over keyframes")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "ms_per_keyframe": 1e3 * dt / args.steps / S, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(B), "keyframes_per_gpu": B, "refs": R, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


NAMES = ["key_img", "key_cam", "T_wk", "ref_imgs", "ref_cams", "T_wr", "age", "prior_depth",
         "prior_variance"]


def run_c5(args, rank, world, dev, barrier):
    """BASELINE.json configs[4] (SURVEY 8d "C5"): n 640x480 keyframe/reference pairs, set S, seeds
    5000 + i, sharded over the ranks in contiguous blocks (tadataka_b200.sharding), every pair its
    own world rendered on its rank by the CUDA ray-caster before the clock starts.  Timed, as one
    region (device events, max over ranks): the kernels of every round of `chunk` pairs plus the
    gather of depth / variance / flag maps on rank 0 (overlapping the next round) plus the
    all-reduce of the flag histogram.  Strong scaling: the total is n pairs at every N."""
    import torch
    import torch.distributed as dist
    from tadataka_b200 import device, scene_device, sharding

    n, chunk = args.c5_pairs, args.c5_chunk
    if world > 1:
        rg = sharding.RootGather(n, H, W, chunk, dev)
    else:
        rg = None
    start, stop = sharding.shard_range(n, rank, world)
    rounds = (max(b - a for a, b in (sharding.shard_range(n, r, world) for r in range(world))) + chunk - 1) // chunk
    inputs = []
    for j in range(rounds):
        a = min(start + j * chunk, stop)
        z = min(a + chunk, stop)
        if z > a:
            t = scene_device.make_keyframes([5000 + i for i in range(a, z)], R=1, H=H, W=W, device=dev,
                                            ages="ones", per_item_rng=True)
            params = np.ascontiguousarray(t["params"], dtype=np.float64)
            inputs.append({k: t[k] for k in NAMES})
        else:
            inputs.append(None)
    if rg is None:
        out = (torch.empty((n, H, W), dtype=torch.float64, device=dev),
               torch.empty((n, H, W), dtype=torch.float64, device=dev),
               torch.empty((n, H, W), dtype=torch.int64, device=dev))
    hist = torch.zeros((chunk, 10), dtype=torch.int64, device=dev)
    params = np.ascontiguousarray(scene_params_s(), dtype=np.float64)

    def one_pass(gather):
        hist.zero_()
        for j in range(rounds):
            t = inputs[j]
            if rg is not None:
                views = rg.out_views(j)
            else:
                a = j * chunk
                views = tuple(o[a:a + (t["key_img"].shape[0] if t else 0)] for o in out)
            if t is not None:
                m = t["key_img"].shape[0]
                device.update_depth(*(t[k] for k in NAMES), params, check=False, out=views, fuse=False,
                                    hist=hist[:m])
            if rg is not None and gather:
                rg.ship(j)
        if rg is not None and gather:
            rg.finish()
        total = hist.sum(dim=0)
        if world > 1:
            dist.all_reduce(total)
        return total

    res = {}
    for label, gather in (("compute_only", False), ("total", True)):
        one_pass(gather)                      # warm-up (allocations, NCCL channels)
        torch.cuda.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        total = one_pass(gather)
        e1.record()
        torch.cuda.synchronize()
        barrier()
        tm = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        res[label] = float(tm[0])
    flags = [int(x) for x in total.tolist()]
    # size-independent checks at the full size: every pixel of every pair carries exactly one flag, and
    # the flag maps that ARRIVED on rank 0 count as many successes as the all-reduced histogram says
    checks = {"histogram_covers_every_pixel": sum(flags) == n * H * W}
    if rank == 0:
        gathered_flag = rg.flag if rg is not None else out[2]
        checks["gathered_flag_maps_match_histogram"] = int((gathered_flag == 0).sum()) == flags[0]
    block = {
        "workload": f"C5: {n} x (640x480 keyframe vs 1 reference frame), sharded over {world} GPU(s) "
                    f"in rounds of {chunk}, every pair its own CUDA-rendered world",
        "pairs": n, "scaling": "strong", "total_ms": res["total"], "ms_per_pair": res["total"] / n,
        "pixels_per_s": n * H * W / (res["total"] * 1e-3),
        "compute_only_ms": res["compute_only"],
        "flag_histogram": flags, "success_px": flags[0], "checks": checks,
    }
    if world > 1:
        rx = rg.bytes_received / 2 if rank == 0 else 0      # two passes shipped (warm-up + timed)
        block["gather"] = {
            "what": "depth f64 + variance f64 + flag int8 of every pair to rank 0 (grouped NCCL send/recv, "
                    "overlapping the next round's kernels) + all-reduce of the 10-bin histogram",
            "bytes_into_rank0": rx, "gbs_into_rank0": rx / (res["total"] * 1e-3) / 1e9,
            "limit": "rank 0's NVLink ingress (900 GB/s nominal per direction)"}
    return block


def scene_params_s():
    from tadataka_b200 import scene
    return scene.SET_S["params"]


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from tadataka_b200 import _lib, device, sharding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; tadataka_b200 has no CPU path "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # one process per GPU: run on (and allocate pinned memory from) the NUMA node next to it
        try:
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        except Exception:
            pass
    if world > 1:
        # NCCL prints a version banner on stdout when the communicator is created; the bench contract
        # wants exactly one JSON line there, so stdout points at stderr while NCCL starts up
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            # NCCL's copy kernels share the SMs with update_depth's grids: on a high-priority stream
            # their CTAs are placed as soon as a compute CTA retires instead of queueing behind the grid
            opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=not args.nccl_default_priority)
            dist.init_process_group("nccl", device_id=dev, pg_options=opts)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            C.CDLL(None).fflush(None)
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    if args.plan >= 0:
        _lib.check(_lib.set_kernel_plan(args.plan))

    B = args.keyframes
    plan = args.plan if args.plan >= 0 else (0 if B * H * W <= 640 * 480 else 2)   # what -1 (automatic) resolves to
    # my kernels per resident step: pair_setup + per sub-batch of <= 64 keyframes (sorted plan)
    # key_prep, sieve, lines, order, match / (two kernels) key_prep, sieve, search / (one kernel) tile
    nsub = (B + 63) // 64
    launches_per_step = {2: 1 + 5 * nsub, 0: 1 + 3 * nsub, 1: 2}[plan]
    kernel_desc = {
        2: "update_depth key_prep + sieve + lines + order + match kernels (step time, incl. the ~us pair_setup_kernel)",
        0: "update_depth key_prep + sieve + search kernels (step time, incl. the ~us pair_setup_kernel)",
        1: "update_depth_tile_kernel (step time, incl. the ~us pair_setup_kernel)"}[plan]
    names = NAMES
    if args.device_scenes:
        # every keyframe its own world, rendered by the CUDA ray-caster (csrc/scene.cu); the host
        # copies below only feed the e2e leg and the CPU baseline
        from tadataka_b200 import scene_device
        t = scene_device.make_keyframes([2000 + 1000 * rank + i for i in range(B)], R=R, H=H, W=W, device=dev)
        b = {k: t[k].cpu().numpy() for k in names}
        b["age"] = b["age"].astype(np.uint64)
        b["params"] = t["params"]
        pinned = {k: torch.from_numpy(b[k]).pin_memory() for k in names}
    else:
        b = make_workload(rank, B, args.distinct)
        pinned = {k: torch.from_numpy(b[k]).pin_memory() for k in names}
        t = {k: pinned[k].to(dev, non_blocking=True) for k in names}
    params = np.ascontiguousarray(b["params"], dtype=np.float64)
    # N > 1: the results of every step are gathered on rank 0 (RootGather, cyclic over the steps)
    rg = sharding.RootGather(world * B, H, W, B, dev, cyclic=True) if world > 1 else None
    out = (torch.empty((B, H, W), dtype=torch.float64, device=dev),
           torch.empty((B, H, W), dtype=torch.float64, device=dev),
           torch.empty((B, H, W), dtype=torch.int64, device=dev))
    hist = torch.zeros((B, 10), dtype=torch.int64, device=dev)
    torch.cuda.synchronize()

    def step(k=0):
        views = rg.out_views(k) if rg is not None else out
        device.update_depth(t["key_img"], t["key_cam"], t["T_wk"], t["ref_imgs"], t["ref_cams"],
                            t["T_wr"], t["age"], t["prior_depth"], t["prior_variance"], params,
                            check=False, out=views, fuse=True, hist=hist if rg is not None else False)
        if rg is not None:
            rg.ship(k)

    def barrier():
        if world > 1:
            dist.barrier()

    nwarm = max(args.warmup, 3)
    for k in range(nwarm):
        step(k)
    if rg is not None:
        rg.finish()
    torch.cuda.synchronize()
    # status of the warm-up launches (age / pose / search-length errors would show here)
    ws = device.workspace(_lib.update_depth_workspace_bytes(B, R, H, W), dev, "update_depth")
    _lib.check(_lib.update_depth_status(C.c_void_p(ws.data_ptr()),
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    sampler = ClockSampler(local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    sampler.start()
    rx0 = rg.bytes_received if rg is not None else 0
    hist.zero_()
    ev0.record()
    for k in range(args.steps):
        step(nwarm + k)
    if rg is not None:
        rg.finish()
        hist_total = hist.sum(dim=0)
        dist.all_reduce(hist_total)
    ev1.record()
    torch.cuda.synchronize()
    sampler.stop_flag = True
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    sampler.join(timeout=1.0)
    gather_info = None
    if rg is not None:
        # the resident maps of this rank's last step (the e2e check below compares with them)
        out = tuple(x.clone() for x in rg.out_views(nwarm + args.steps - 1)) if rank != 0 else \
            tuple(x[:B].clone() for x in (rg.depth, rg.variance, rg.flag))
        rx = rg.bytes_received - rx0
        gather_info = {"bytes_into_rank0_per_step": rx / max(args.steps, 1),
                       "what": "depth f64 + variance f64 + flag int8 of every keyframe of every step to rank 0 "
                               "(grouped NCCL send/recv overlapping the next step) + one all-reduce of the "
                               "10-bin histograms", "inside_timed_region": True}
        # beside it, the same loop with less on the wire: the depth maps only (what BASELINE.json's
        # configs[4] names) and nothing at all (the kernels alone) -- what the link costs
        for label, maps in (("depth_maps_only", ("depth",)), ("no_gather", None)):
            rg2 = sharding.RootGather(world * B, H, W, B, dev, cyclic=True, maps=maps) if maps else None
            views2 = out if rg2 is None else None
            for k in range(3 + args.steps):
                if k == 3:
                    if rg2 is not None:
                        rg2.finish()
                    torch.cuda.synchronize()
                    barrier()
                    ev0.record()
                v = rg2.out_views(k) if rg2 is not None else views2
                device.update_depth(t["key_img"], t["key_cam"], t["T_wk"], t["ref_imgs"], t["ref_cams"],
                                    t["T_wr"], t["age"], t["prior_depth"], t["prior_variance"], params,
                                    check=False, out=v, fuse=True, hist=hist)
                if rg2 is not None:
                    rg2.ship(k)
            if rg2 is not None:
                rg2.finish()
            h2 = hist.sum(dim=0)
            dist.all_reduce(h2)
            ev1.record()
            torch.cuda.synchronize()
            barrier()
            tm2 = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
            dist.all_reduce(tm2, op=dist.ReduceOp.MAX)
            gather_info[label] = {"ms_per_step": float(tm2[0]) / args.steps,
                                  "value": world * B * H * W * args.steps / (float(tm2[0]) * 1e-3)}
            del rg2

    # ---- latency of ONE keyframe (8 references) per launch: "ms per keyframe-pair" without batching
    one = {k: t[k][:1].contiguous() for k in names}
    one_out = tuple(o[:1].contiguous() for o in out)
    lat0, lat1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        device.update_depth(*(one[k] for k in names), params, check=False, out=one_out, fuse=True)
    torch.cuda.synchronize()
    lat0.record()
    for _ in range(50):
        device.update_depth(*(one[k] for k in names), params, check=False, out=one_out, fuse=True)
    lat1.record()
    torch.cuda.synchronize()
    single_ms = lat0.elapsed_time(lat1) / 50

    # ---- secondary shape (BASELINE.json configs[3] "C4"): 64 independent 1280x720 TUM-scale pairs
    # per launch (every pair its own CUDA-rendered world), kernels only
    c4_ms = None
    if rank == 0 and not args.no_c4 and args.c4_pairs > 0:
        from tadataka_b200 import scene, scene_device
        n4 = args.c4_pairs
        c4t = scene_device.make_keyframes([4000 + i for i in range(n4)], R=1, H=720, W=1280, cam=scene.CAM_720P,
                                          pset=scene.SET_T, device=dev, ages="ones", per_item_rng=True)
        c4p = np.ascontiguousarray(c4t["params"], dtype=np.float64)
        c4o = tuple(torch.empty((n4, 720, 1280), dtype=dt, device=dev)
                    for dt in (torch.float64, torch.float64, torch.int64))
        for _ in range(3):
            device.update_depth(*(c4t[k] for k in names), c4p, check=False, fuse=True, out=c4o)
        torch.cuda.synchronize()
        lat0.record()
        for _ in range(10):
            device.update_depth(*(c4t[k] for k in names), c4p, check=False, fuse=True, out=c4o)
        lat1.record()
        torch.cuda.synchronize()
        c4_ms = lat0.elapsed_time(lat1) / 10 / n4
        del c4t, c4o
        torch.cuda.empty_cache()

    # ---- secondary loop (BASELINE.json configs[2] "C3"): the semi_dense_vo keyframe loop over 30
    # frames (increment_age -> propagate -> update_depth + fusion -> regularize, unused reference
    # frames dropped), maps and reference frames device resident, one image upload per frame;
    # wall clock per frame.  Frames come from the CUDA ray-caster (host copies, as a caller has them).
    c3_ms = c3_f64_ms = None
    if rank == 0 and not args.no_c4:
        from rust_bindings.camera import CameraParameters
        from tadataka_b200 import scene, scene_device
        from tadataka_b200.mapper import SemiDenseMapper
        seq_poses = scene.sequence_poses(3300, 30)
        seq_world = scene.World(scene.SET_S, 3300)
        seq_img, _ = scene_device.render(seq_world, seq_poses, np.tile(scene.CAM_VGA, (30, 1)), H, W, dev)
        seq_img = seq_img.cpu().numpy()
        cp = CameraParameters((scene.CAM_VGA[0], scene.CAM_VGA[1]), (scene.CAM_VGA[2], scene.CAM_VGA[3]))
        depth0 = np.random.default_rng(3300).uniform(params[0], params[1], (H, W))
        # a camera delivers 8-bit frames: the loop is timed on those (converted to the f64 gray image on
        # the device, 0.3 MB per upload) and, for comparison, on f64 frames (2.4 MB per upload)
        seq_u8 = np.round(seq_img * 255.0).astype(np.uint8)
        c3 = {}
        for label, seq in (("f64", seq_img), ("u8", seq_u8)):
            for rep in range(2):   # the first pass warms allocator and workspaces
                mapper = SemiDenseMapper(cp, seq[0], seq_poses[0], params, 200.0, 100.0, 1e-6, depth_map=depth0,
                                         fuse=True, regularize=True, prune=True)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for i in range(1, 30):
                    mapper.track(cp, seq[i], seq_poses[i])
                mapper.sync()
                c3[label] = (time.perf_counter() - t0) * 1e3 / 29
        c3_ms, c3_f64_ms = c3["u8"], c3["f64"]
        del mapper

    # ---- the companion kernels on 64 VGA frames (increment_age, propagate, fusion, regularize)
    companions = None
    if rank == 0 and world == 1 and not args.no_c4:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import companion_bench
        companions = companion_bench.run(frames=64, iters=10)
        torch.cuda.empty_cache()
    # ---- the numpy drop-in, one call per frame (what a user of the reference's module writes)
    dropin = None
    if rank == 0 and world == 1 and not args.no_c4:
        import dropin_bench
        dropin = dropin_bench.run(refs=R, iters=20)
        # the tracker on the other side of the loop: the 5-level direct pose-change
        # estimator on one VGA pair, numpy images in, pyramid and Gauss-Newton loops on the device
        from rust_bindings.camera import CameraParameters
        from tadataka_b200 import dvo, scene as _scene
        _w = _scene.World(_scene.SET_S, 9100)
        _cam = _scene.CAM_VGA
        _I0, _D0 = _w.render(_scene.pose(), _cam, H, W)
        _I1, _ = _w.render(_scene.pose(_scene.rotation(0.012, -0.018, 0.009), (6.0, -2.5, 3.5)), _cam, H, W)
        _cp = CameraParameters((_cam[0], _cam[1]), (_cam[2], _cam[3]))
        _est = dvo.PoseChangeEstimator(_cp, _cp, n_coarse_to_fine=5, max_iter=20)
        _est(_I0, _D0, _I1)
        torch.cuda.synchronize()
        _t0 = time.perf_counter()
        for _ in range(5):
            _est(_I0, _D0, _I1)
        dropin["dvo_pose_change_5_levels_ms"] = 1e3 * (time.perf_counter() - _t0) / 5

    # ---- C5: the sharded run with the gather inside the clock (all ranks take part)
    c5 = None
    if args.c5_pairs > 0:
        c5 = run_c5(args, rank, world, dev, barrier)
        torch.cuda.empty_cache()

    # ---- e2e: host-pointer C-ABI call, pinned host buffers, copies inside the timed region
    e2e_steps = args.e2e_steps or min(args.steps, 30)
    h_out = (torch.empty((B, H, W), dtype=torch.float64).pin_memory(),
             torch.empty((B, H, W), dtype=torch.float64).pin_memory(),
             torch.empty((B, H, W), dtype=torch.int64).pin_memory())

    def vp(x):
        return C.c_void_p(x.data_ptr())

    # frames by pointer: keyframe k of the step shows scene k % distinct, so its key image and its 8
    # reference images are the host arrays of that scene -- an image several keyframes name crosses
    # PCIe once per call (tdk_update_depth_frames_host), as in the reference's loop, where every
    # keyframe is matched against frames that are already there
    nd = B if args.device_scenes else max(1, min(args.distinct, B))
    hwb = H * W * 8
    key_ptrs = (C.c_void_p * B)(*[pinned["key_img"].data_ptr() + (k % nd) * hwb for k in range(B)])
    ref_ptrs = (C.c_void_p * (B * R))(*[pinned["ref_imgs"].data_ptr() + ((k % nd) * R + r) * hwb
                                        for k in range(B) for r in range(R)])

    def e2e_step():
        rc = _lib.update_depth_frames_host(key_ptrs, vp(pinned["key_cam"]), vp(pinned["T_wk"]), ref_ptrs,
                                           vp(pinned["ref_cams"]), vp(pinned["T_wr"]), vp(pinned["age"]),
                                           vp(pinned["prior_depth"]), vp(pinned["prior_variance"]),
                                           params.ctypes.data_as(C.c_void_p), B, R, H, W, vp(h_out[0]), vp(h_out[1]),
                                           vp(h_out[2]), None, None, 1)
        _lib.check(rc)

    def e2e_step_private():
        rc = _lib.update_depth_host(vp(pinned["key_img"]), vp(pinned["key_cam"]), vp(pinned["T_wk"]),
                                    vp(pinned["ref_imgs"]), vp(pinned["ref_cams"]), vp(pinned["T_wr"]),
                                    vp(pinned["age"]), vp(pinned["prior_depth"]),
                                    vp(pinned["prior_variance"]), params.ctypes.data_as(C.c_void_p),
                                    B, R, H, W, vp(h_out[0]), vp(h_out[1]), vp(h_out[2]), None, None, 1)
        _lib.check(rc)

    # secondary: every keyframe ships private copies of its 8 reference images (tdk_update_depth_host)
    e2e_step_private()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(max(1, e2e_steps // 2)):
        e2e_step_private()
    torch.cuda.synchronize()
    e2e_private_s = (time.perf_counter() - t0) / max(1, e2e_steps // 2)
    h2d_private = sum(pinned[k].numel() * pinned[k].element_size() for k in names)

    e2e_step()
    e2e_step()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    small = sum(pinned[k].numel() * pinned[k].element_size() for k in ("key_cam", "T_wk", "ref_cams", "T_wr"))
    # bytes that cross the link: images and f64 maps as they are; the u64 ages and i64 flags of the
    # caller's arrays travel as one byte per pixel (narrowed / widened by the library's host workers)
    h2d = nd * (1 + R) * hwb + 2 * B * hwb + B * H * W + small
    d2h = 2 * B * hwb + B * H * W
    # the e2e path must give the same answer as the resident path
    same = bool(torch.equal(h_out[0], out[0].cpu()) and torch.equal(h_out[1], out[1].cpu())
                and torch.equal(h_out[2], out[2].cpu()))

    # ---- aggregate over ranks (max time)
    if world > 1:
        tm = torch.tensor([ms_total, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms_total, e2e_s = float(tm[0]), float(tm[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    px_step = B * H * W
    ms_step = ms_total / args.steps
    value = world * px_step * args.steps / (ms_total * 1e-3)
    peak, peak_src = measured_peak()
    achieved = ALGO_BYTES_PER_PX * px_step / (ms_step * 1e-3) / 1e9
    traffic_px = profiled_traffic()
    n_success = int((out[2] == 0).sum())
    if gather_info is not None:
        gather_info["gbs_into_rank0"] = gather_info["bytes_into_rank0_per_step"] / (ms_step * 1e-3) / 1e9
        gather_info["limit"] = "rank 0's NVLink ingress (900 GB/s nominal per direction)"
        gather_info["success_px_all_ranks"] = int(hist_total[0])
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": nwarm, "ms_per_step": ms_step, "ms_per_keyframe": ms_step / B,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {
            "workload": workload_name(B),
            "keyframes_per_gpu": B, "refs": R,
            "distinct_scenes": B if args.device_scenes else min(args.distinct, B),
            "scenes": "CUDA ray-caster" if args.device_scenes else "numpy ray-caster",
            "l2": f"inputs {h2d_private / 1e9:.2f} GB per step >> 126 MB L2 (no flush needed)",
            "success_px_per_step": n_success, "success_px_per_s": world * n_success / (ms_step * 1e-3),
            "params": [float(x) for x in params], "e2e_matches_resident": same,
            "nccl_gather": gather_info, "single_keyframe_launch_ms": single_ms,
            "c4_720p_ms_per_pair": c4_ms, "c4_pairs_per_launch": args.c4_pairs if c4_ms is not None else None,
            "c3_vo_loop_ms_per_frame": c3_ms, "c3_vo_loop_f64_frames_ms_per_frame": c3_f64_ms,
            "c3_frames": "30 x 640x480, 8-bit gray frames from the host, one tdk_mapper_track call per frame",
            "companions": companions, "dropin_ms_per_call": dropin, "c5": c5,
        },
        "roofline": {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "peak_source": peak_src,
            "traffic": None if traffic_px is None else traffic_px * px_step,
            "kernel": kernel_desc,
            "algorithmic_bytes_per_px": ALGO_BYTES_PER_PX,
        },
        "clocks": sampler.result(),
        "e2e": {"value": world * px_step * e2e_steps / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "ms_per_step": 1e3 * e2e_s / e2e_steps,
                "api": "tdk_update_depth_frames_host (pinned host buffers, u64 ages and i64 flags cross the link as bytes, frames by pointer: the step's "
                       f"{nd * (1 + R)} distinct images cross PCIe once, fuse_on_success=1)",
                "private_copies": {"value": world * px_step / e2e_private_s, "ms_per_step": 1e3 * e2e_private_s,
                                   "h2d_bytes_per_step": h2d_private,
                                   "api": "tdk_update_depth_host ([P,R,H,W]: every keyframe ships its own 8 reference images)"}},
        "gpu_launches": launches_per_step * args.steps,
    }
    if world == 1 and not args.no_cpu_baseline:
        import oracle as ora
        threads = host_threads()
        n_kf = B
        cpu_maps = []
        v, best, reps = time_cpu(ora, b, n_kf, threads, keep=cpu_maps)
        v1, _, _ = time_cpu(ora, b, 2, 1, budget_s=3.0, max_reps=3)
        # the CPU run and the GPU run did the same work: identical maps on every keyframe of the step
        same_maps = all(np.array_equal(c, g[:n_kf].cpu().numpy(), equal_nan=True)
                        for c, g in zip(cpu_maps, out))
        line["cpu_baseline"] = {
            "value": v, "unit": UNIT, "cores": threads, "kind": "port", "single_thread_value": v1,
            "maps_identical_to_gpu": bool(same_maps),
            "sample": f"all {n_kf} keyframes of the step, best of {reps} runs, "
                      f"{1e3 * best / n_kf:.2f} ms/keyframe, OpenMP over keyframes"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def main():
    args = parse()
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
               str(free_port()), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        # the host-pointer entry point converts ages and flags on a pool of host threads: share the cores between ranks
        os.environ.setdefault("TDK_HOST_THREADS", str(max(2, min(6, (os.cpu_count() or 8) // max(1, world)))))
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
