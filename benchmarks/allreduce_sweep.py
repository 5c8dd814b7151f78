"""Bucket all-reduce bandwidth sweep, 1 KB - 1 GB: our P2P kernels vs NCCL (BASELINE.json config 5).

    torchrun --nproc-per-node N benchmarks/allreduce_sweep.py [--max-mb 1024] [--dtype fp32|bf16]

Times every size with CUDA events on the launching stream (max over ranks), reports algorithm
bandwidth S/t and bus bandwidth S/t * 2(N-1)/N (nccl-tests definition), and the fraction of the
measured NVLink peer-copy roofline (770 GB/s per direction, B200_PROFILING.md).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import distributeddataparallel_b200 as d
from distributeddataparallel_b200.parallel.p2p import P2PComm

p = argparse.ArgumentParser()
p.add_argument("--max-mb", type=int, default=1024)
p.add_argument("--dtype", default="fp32")
p.add_argument("--iters", type=int, default=20)
p.add_argument("--blocks", type=int, default=None)
p.add_argument("--out", default=None)
p.add_argument("--sizes", default="all", help="all | buckets (the dpp.py bucket sizes + 1 MiB, 32 MiB, 256 MiB)")
p.add_argument("--no-nccl", action="store_true")
p.add_argument("--graph", action="store_true", help="time 10 back-to-back launches replayed from a CUDA graph "
               "(device time per collective, no host launch overhead)")
a = p.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
torch.cuda.set_device(dev)
d.setup(rank, world, backend="nccl")
dtype = torch.float32 if a.dtype == "fp32" else torch.bfloat16
esize = 4 if dtype == torch.float32 else 2
comm = P2PComm(None, dev, max_blocks=a.blocks)
comm._arena_bytes = (a.max_mb << 20) + (32 << 20)
comm.reserve(0)
LINK = 770.0   # GB/s per direction per GPU, measured peer copy (B200_PROFILING.md)


def timed(fn, iters):
    if a.graph:
        return timed_graph(fn, iters)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())     # ms, max over ranks


def timed_graph(fn, iters, per_graph=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(per_graph):
            fn()
    g.replay()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    reps = max(1, iters // per_graph)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (reps * per_graph)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


big = comm.allocate((a.max_mb << 20) // esize, dtype, dev)
rows = []
sizes = [1 << k for k in range(10, 31) if (1 << k) <= (a.max_mb << 20)]
sizes += [9_461_800 // 16 * 16, 26_494_976, 8_769_792, 44_726_568 // 16 * 16, 38_144, 921_600, 7_810_048]     # the dpp.py buckets / launch units
if a.sizes == "buckets":
    sizes = [38_144, 921_600, 1 << 20, 7_810_048, 8_769_792, 9_461_792, 26_494_976, 32 << 20, 256 << 20]
    sizes = [s for s in sizes if s <= (a.max_mb << 20)]
for nbytes in sorted(set(sizes)):
    flat = big[: nbytes // esize]
    flat.fill_(1.0)
    ref = torch.ones(nbytes // esize, dtype=dtype, device=dev)
    iters = a.iters if nbytes <= (64 << 20) else max(5, a.iters // 4)
    res = {"bytes": nbytes}
    algos = ("push", "oneshot", "twoshot") + (("nvls",) if comm.arena is not None and comm.arena.multicast else ())
    for algo in algos:
        if algo == "oneshot" and nbytes > (4 << 20):
            continue
        if algo == "push" and nbytes > comm.push_max:
            continue
        comm.algo = algo
        ms = timed(lambda: comm.launch(0, flat, True).wait(), iters)
        res[algo + "_us"] = ms * 1e3
    best = min(v for k, v in res.items() if k.endswith("_us"))
    comm.algo = "auto"
    res["auto_picks"] = comm.ALGO_NAMES[comm._pick_algo(nbytes)]
    nccl_ms = float("nan") if a.no_nccl else timed(lambda: dist.all_reduce(ref), iters)
    res["nccl_us"] = nccl_ms * 1e3
    f = 2 * (world - 1) / world
    res["ours_busbw_GBs"] = nbytes / (best * 1e-6) * f / 1e9
    res["nccl_busbw_GBs"] = nbytes / (nccl_ms * 1e-3) * f / 1e9
    # per GPU and direction: (N-1)/N*S of loads + (N-1)/N*S of pushes  ->  t >= 2(N-1)/N * S / link
    res["roofline_us"] = 2 * (world - 1) / world * nbytes / (LINK * 1e9) * 1e6
    res["frac_of_nvlink_roofline"] = res["roofline_us"] / best
    rows.append(res)
    if rank == 0:
        print(json.dumps(res), flush=True)
comm.check()
if rank == 0 and a.out:
    with open(a.out, "w") as fh:
        json.dump({"world": world, "dtype": a.dtype, "link_GBs": LINK, "graph": bool(a.graph),
                   "arena": comm.arena.provider, "multicast": bool(comm.arena.multicast),
                   "nvls_blocks": comm.nvls_blocks, "blocks": comm.max_blocks, "rows": rows}, fh, indent=1)
d.cleanup()
