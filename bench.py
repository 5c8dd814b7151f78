#!/usr/bin/env python
"""Headline benchmark: samples/s of the dpp.py training step (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference]

Config (both arms): the reference's model and loop -- SimpleCNN = ResNet-18 with a 10-way
head, fp32 parameters, batch 32 per rank, CrossEntropy, SGD lr 0.01, DDP over N ranks
(/root/reference/dpp.py:11-18, 35, 38-41) -- on synthetic CIFAR-10-shaped data with
random-init weights (no network on the box).  Weak scaling: per-GPU batch is fixed.

Arms
  ours       this repo: sm_100a kernels, C++ reducer, P2P all-reduce, whole step in a CUDA graph.
  reference  the UNMODIFIED dpp.py installed under baseline/_ref, driven through its own
             ``train(rank, world_size)``; only its network-dependent imports are shimmed
             (CIFAR-10 download -> synthetic arrays with torchvision's own ``__getitem__``;
             ``pretrained=True`` -> random init) and its ``DataLoader`` is wrapped to place the
             timing marks.  None of this repo's models/kernels/engine are on that path.

Numbers
  value      device-timed (CUDA events, max over ranks) samples/s over K steps whose inputs are
             already on the device; inputs rotate through a pool larger than L2.
  e2e.value  the same K steps through the public API including, per step, the host->device
             copy of the batch from pinned memory and a device->host read of the loss.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 32               # dpp.py:35
L2_BYTES = 126 * 2**20


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--comm", default="auto", help="ours: auto|p2p|c10d")
    p.add_argument("--algo", default="auto")
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--no-native", action="store_true", help="ours: library ops instead of the sm_100a kernels")
    p.add_argument("--model", default="simplecnn",
                   help="simplecnn (headline, BASELINE config 2) | resnet50 (config 3) | gpt2-medium (config 4)")
    p.add_argument("--batch", type=int, default=None, help="per-GPU batch (32; 8 sequences for gpt2)")
    p.add_argument("--dtype", default=None, help="extra workloads: bf16 | tf32 (default: bf16 for gpt2, tf32 for resnet50)")
    p.add_argument("--seq-len", type=int, default=1024)
    return p.parse_args()


def dist_env(args):
    """(rank, local_rank, world) -- re-exec under torchrun when N>1 was asked for directly."""
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        return int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"])), int(os.environ["WORLD_SIZE"])
    if args.gpus > 1:
        import socket

        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    return 0, 0, 1


def max_over_ranks(x: float, world: int, device) -> float:
    import torch
    import torch.distributed as dist

    if world == 1 or not dist.is_initialized():
        return x
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    import torch
    import torch.distributed as dist

    if world > 1 and dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize()


def load_clock_sampler():
    """``ClockSampler`` loaded by file path: importing the package would map this repository's native
    libraries into the process, which the reference arm must never do."""
    import importlib.util

    path = os.path.join(ROOT, "distributeddataparallel_b200", "utils", "clocks.py")
    spec = importlib.util.spec_from_file_location("_b200ddp_clocks", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ClockSampler


def find_reference(rank: int, world: int):
    """Directory holding the unmodified ``dpp.py``: ``baseline/_ref`` (written by ``build()``), else installed
    now from ``$B200DDP_REFERENCE`` / ``/root/reference`` (byte-identical copy), else None."""
    import shutil

    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    dst = os.path.join(ref_dir, "dpp.py")
    if os.path.isfile(dst):
        return ref_dir
    cands = [os.environ.get("B200DDP_REFERENCE"), "/root/reference"]
    for c in cands:
        if not c:
            continue
        src = os.path.join(c, "dpp.py") if os.path.isdir(c) else c
        if not os.path.isfile(src):
            continue
        try:
            if int(os.environ.get("LOCAL_RANK", rank)) == 0:
                os.makedirs(ref_dir, exist_ok=True)
                tmp = dst + f".tmp{os.getpid()}"
                shutil.copyfile(src, tmp)
                os.replace(tmp, dst)
            else:
                for _ in range(200):
                    if os.path.isfile(dst):
                        break
                    time.sleep(0.05)
            if os.path.isfile(dst) and open(dst, "rb").read() == open(src, "rb").read():
                return ref_dir
        except OSError:
            pass
        return os.path.dirname(os.path.abspath(src))       # read-only tree: import it from where it is
    return None


def loaded_native_libs():
    """Shared objects from this repository mapped into the current process (the reference arm must show [])."""
    out = set()
    try:
        with open("/proc/self/maps") as f:
            for line in f:
                path = line.split()[-1] if line.strip() else ""
                if path.startswith(ROOT) and ".so" in path and "/baseline/" not in path:
                    out.add(os.path.relpath(path, ROOT))
    except OSError:
        pass
    return sorted(out)


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner, the reference's loss prints) write to fd 1; the contract is
    ONE JSON line on stdout, so everything else is routed to stderr for the duration of the run."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(rank, payload):
    if rank == 0:
        out = _REAL_STDOUT or sys.stdout
        out.write(json.dumps(payload) + "\n")
        out.flush()


# =========================================================================================
# our arm
# =========================================================================================
def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    import distributeddataparallel_b200 as d

    ClockSampler = load_clock_sampler()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        d.setup(rank, world, backend="nccl")
    d.ops.set_native(not args.no_native)
    torch.manual_seed(0)
    cfg = d.TrainConfig(model=args.model, batch_size=args.batch)
    model = d.build_model(cfg).to(dev)
    ddp = d.DDP(model, device_ids=[local_rank], comm=args.comm, allreduce_algo=args.algo)
    crit = d.CrossEntropyLoss()
    opt = d.SGD(ddp.parameters(), lr=0.01)
    step = d.TrainStep(ddp, crit, opt, cuda_graph=not args.no_graph)
    ddp.train()
    K, W, B = args.steps, args.warmup, args.batch

    # ---- device-resident input pool larger than L2 (rotated every step) -------------------
    per_batch = B * 3 * 32 * 32 * 4
    pool_n = max(8, -(-int(1.25 * L2_BYTES) // per_batch))
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    pool_u8 = torch.randint(0, 256, (pool_n, B, 32, 32, 3), dtype=torch.uint8, generator=g)
    pool_y = torch.randint(0, 10, (pool_n, B), dtype=torch.int64, generator=g).to(dev)
    xs = [d.ops.preprocess_u8(pool_u8[i].to(dev)) for i in range(pool_n)]     # NHWC fp32, one allocation each

    native = None
    if not args.no_native:
        from distributeddataparallel_b200.ops import native as native_mod

        native = native_mod

    # ---- internal settle (bucket rebuild, graph capture) + W warm-up steps ------------------
    sampler = ClockSampler(gpu_index=local_rank).start()
    for i in range(4 + W):
        step(xs[i % pool_n], pool_y[i % pool_n])
    sampler.wait_for_samples(1)
    if K * 1.5 < 300:                      # short run: keep the load on until the sampler has something to show
        for i in range(64):
            step(xs[i % pool_n], pool_y[i % pool_n])
    barrier(world)
    launches_before = native.launch_count() if native else 0
    replays_before = step.replays
    clock_mark = sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        loss = step(xs[i % pool_n], pool_y[i % pool_n])
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    dev_ms = max_over_ranks(e0.elapsed_time(e1), world, dev)
    clock_end = sampler.mark()
    clocks = sampler.summary(clock_mark, clock_end)
    launches = 0
    if native:
        eager_launches = native.launch_count() - launches_before
        launches = eager_launches + (step.replays - replays_before) * native.launches_per_capture()
    loss_val = float(loss.item())

    # ---- end to end through the public API: pinned H2D + step + D2H loss, every step --------
    ds = d.SyntheticCIFAR10(n=(2 * K + W + 8) * B * world, seed=99)
    samp = d.DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True)
    loader = d.DeviceLoader(ds, B, samp, dev)
    it = iter(loader)
    for _ in range(W):
        x, y = next(it)
        step(x, y).item()
    # (a) every step's loss is read back with the framework's AsyncLossReader: a pinned D2H copy right behind each
    #     step, consumed by the host one step later, so the read-back does not drain the GPU pipeline
    reader = d.AsyncLossReader()
    barrier(world)
    t0 = time.perf_counter()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    host_loss = 0.0
    for _ in range(K):
        x, y = next(it)                       # pinned uint8 batch -> H2D -> fused normalise
        reader.push(step(x, y))               # D2H read of the loss, every step (async, pinned)
        if reader.pending() > 1:
            host_loss = reader.pop()
    for v in reader.drain():
        host_loss = v
    c1.record()
    torch.cuda.synchronize()
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max_over_ranks(max(c0.elapsed_time(c1), e2e_wall_ms), world, dev)
    barrier(world)
    # (b) the same loop with a blocking loss.item() per step (a full device synchronisation every step)
    t0 = time.perf_counter()
    for _ in range(K):
        x, y = next(it)
        host_loss_blocking = step(x, y).item()
    torch.cuda.synchronize()
    e2e_blocking_ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world, dev)
    barrier(world)

    sampler.stop()
    samples = K * B * world
    value = samples / (dev_ms / 1e3)
    log = ddp._get_ddp_logging_data()
    # ---- the other half of the metric: device-timed bucket all-reduce at this N -------------------------
    bucket_sizes = log.get("rebuilt_bucket_sizes") or log.get("initial_bucket_sizes") or []
    allreduce = []
    timer = getattr(ddp.comm, "time_allreduce", None)
    if world > 1 and timer is not None:
        for nbytes in bucket_sizes:
            allreduce.append(timer(int(nbytes), iters=20))
    # ---- diagnostics (outside every timed region): device-stamped breakdown of one replayed step ----------------
    breakdown = None
    if not args.no_graph and hasattr(ddp, "enable_device_timing"):
        try:
            ddp.enable_device_timing()
            step2 = d.TrainStep(ddp, crit, opt, cuda_graph=True)          # re-captured WITH the time-stamp marks
            for i in range(8):
                step2(xs[i % pool_n], pool_y[i % pool_n])
            breakdown = ddp.device_timing()
            exposed = max_over_ranks(float(breakdown.get("exposed_comm_us", 0.0)), world, dev)
            breakdown["exposed_comm_us_max_over_ranks"] = exposed
            if breakdown.get("step_us"):
                breakdown["exposed_comm_pct_max_over_ranks"] = 100.0 * exposed / breakdown["step_us"]
            breakdown["note"] = "one replayed step re-captured with 4 stamp kernels (adds a few us); not the timed run"
            del step2
        except Exception as e:  # diagnostics must never cost the headline
            breakdown = {"error": repr(e)[:200]}
    payload = {
        "metric": "samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32 storage, tf32 tensor-core math (reference: fp32 with cuDNN TF32 convs)",
        "data": "synthetic CIFAR-10-shaped uint8 images, random-init weights",
        "impl": "ours",
        "config": {"model": "dpp.py SimpleCNN (ResNet-18, 10 classes)", "global_batch": B * world, "per_gpu_batch": B,
                   "image": "3x32x32", "optimizer": "SGD lr=0.01", "loss": "cross-entropy",
                   "parallelism": f"dp{world}", "comm": ddp.comm.name, "cuda_graph": step.captured,
                   "native_kernels": not args.no_native,
                   "l2_policy": f"inputs rotate through a {pool_n * per_batch / 2**20:.0f} MiB device pool (> 126 MiB L2); "
                                "model state (params+grads+buckets ~134 MB) also exceeds L2",
                   "buckets": log.get("rebuilt_bucket_sizes") or log.get("initial_bucket_sizes")},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"),
                   "reasons": clocks.get("reasons", []), "samples_in_region": 0 if clocks.get("pre_region_fallback")
                   else clocks.get("samples", 0), "pre_region_fallback": bool(clocks.get("pre_region_fallback")),
                   "source": clocks.get("source")},
        "allreduce": allreduce,
        "step_breakdown": breakdown,
        "comm_stats": log.get("comm_stats"),
        "e2e": {"value": samples / (e2e_ms / 1e3), "unit": "samples/s", "ms_per_step": e2e_ms / K,
                "h2d_bytes_per_step": loader.h2d_bytes_per_batch, "d2h_bytes_per_step": 4,
                "api": "DeviceLoader -> TrainStep(DDP(SimpleCNN), CrossEntropyLoss, SGD) -> AsyncLossReader "
                       "(pinned D2H of every step's loss, consumed by the host one step later)",
                "blocking_item": {"value": samples / (e2e_blocking_ms / 1e3), "ms_per_step": e2e_blocking_ms / K,
                                  "api": "... -> loss.item() every step (device synchronisation per step)"}},
        "gpu_launches": int(launches),
        "final_loss": loss_val, "e2e_final_loss": host_loss,
        "graph_replays": step.replays, "grads_in_place": log.get("grads_produced_in_place"),
        "grads_copied_in": log.get("grads_copied_in"),
    }
    emit(rank, payload)
    del step
    if world > 1:
        d.cleanup()


# =========================================================================================
# reference arm: unmodified dpp.py
# =========================================================================================
def run_reference(args, rank, local_rank, world):
    ref_dir = find_reference(rank, world)
    if ref_dir is None:
        emit(rank, {"impl": "reference", "unavailable": "dpp.py not found in baseline/_ref, $B200DDP_REFERENCE or "
                                                        "/root/reference (run __graft_entry__.build() first)"})
        return
    try:
        import numpy as np
        import torch
        import torch.distributed as dist
        import torchvision
        from torch.utils.data import DataLoader as StockLoader
    except Exception as e:  # pragma: no cover
        emit(rank, {"impl": "reference", "unavailable": f"import failed: {e!r}"[:200]})
        return
    sys.path.insert(0, ref_dir)
    import dpp  # the unmodified reference module

    ClockSampler = load_clock_sampler()      # by path: none of this repository's native code is mapped here

    K, W, B = args.steps, args.warmup, BATCH
    need = (2 * (K + W) + 8) * B * world

    class SynthCIFAR10(torchvision.datasets.CIFAR10):
        """torchvision's CIFAR10 (stock __getitem__: PIL image -> transform) on random pixels."""

        def __init__(self, root, train=True, transform=None, target_transform=None, download=False):
            torchvision.datasets.VisionDataset.__init__(self, root, transform=transform,
                                                        target_transform=target_transform)
            rng = np.random.default_rng(1234)
            self.train = train
            self.data = rng.integers(0, 256, size=(need, 32, 32, 3), dtype=np.uint8)
            self.targets = rng.integers(0, 10, size=(need,)).tolist()
            self.classes = [str(i) for i in range(10)]

    class _Datasets:
        CIFAR10 = SynthCIFAR10

    class _Models:
        @staticmethod
        def resnet18(pretrained=False, **kw):          # no network: random init
            return torchvision.models.resnet18(weights=None)

    marks = {}

    class TimedLoader:
        """Wraps the stock DataLoader the reference constructs; places timing marks.

        phase A  W warm-up + K timed steps on batches collated beforehand (no host transform
                 work between steps)  -> "value"
        phase B  W warm-up + K timed steps straight from the stock DataLoader (PIL + ToTensor +
                 Normalize per sample, num_workers=0, as dpp.py:32-35 does)  -> "e2e"
        """

        def __init__(self, dataset, batch_size=1, sampler=None, **kw):
            self.inner = StockLoader(dataset, batch_size=batch_size, sampler=sampler, **kw)

        def __len__(self):
            return 2 * (K + W)

        def _sync(self):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        def __iter__(self):
            it = iter(self.inner)
            ready = [next(it) for _ in range(K + W)]
            marks["clock"] = ClockSampler(gpu_index=local_rank).start()
            for i, b in enumerate(ready):
                if i == W:
                    self._sync()
                    marks["clock"].wait_for_samples(1)
                    marks["clock_mark"] = marks["clock"].mark()
                    marks["a0"] = torch.cuda.Event(enable_timing=True)
                    marks["a0"].record()
                yield b
            marks["a1"] = torch.cuda.Event(enable_timing=True)
            marks["a1"].record()
            torch.cuda.synchronize()
            marks["clocks"] = marks["clock"].stop(marks.get("clock_mark", 0))
            marks["a_loaded"] = loaded_native_libs()
            for i in range(K + W):
                if i == W:
                    self._sync()
                    marks["b_t0"] = time.perf_counter()
                yield next(it)
            torch.cuda.synchronize()
            marks["b_t1"] = time.perf_counter()
            # the reference's process group is still up here (dpp.cleanup runs after the loop):
            # reduce the two timings to their max over ranks now
            a_ms = marks["a0"].elapsed_time(marks["a1"])
            b_ms = (marks["b_t1"] - marks["b_t0"]) * 1e3
            if world > 1:
                t = torch.tensor([a_ms, b_ms], dtype=torch.float64, device=torch.device("cuda", local_rank))
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                a_ms, b_ms = float(t[0]), float(t[1])
            marks["a_ms"], marks["b_ms"] = a_ms, b_ms

    dpp.datasets = _Datasets
    dpp.models = _Models
    dpp.DataLoader = TimedLoader
    torch.cuda.set_device(local_rank)
    with contextlib.redirect_stdout(sys.stderr):
        dpp.train(rank, world, epochs=1)          # the reference's own public per-rank API
    a_ms, b_ms = marks["a_ms"], marks["b_ms"]
    samples = K * B * world
    clocks = marks.get("clocks", {})
    emit(rank, {
        "metric": "samples/sec", "value": samples / (a_ms / 1e3), "unit": "samples/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": a_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32 (cuDNN TF32 convs allowed, stock defaults)",
        "data": "synthetic CIFAR-10-shaped uint8 images, random-init weights", "impl": "reference",
        "config": {"model": "dpp.py SimpleCNN (ResNet-18, 10 classes)", "global_batch": B * world,
                   "per_gpu_batch": B, "parallelism": f"dp{world}", "comm": "nccl (torch DDP)",
                   "l2_policy": "fresh batch copied from host every step; model state ~134 MB exceeds L2",
                   "value_definition": "unmodified dpp.train loop fed batches collated beforehand"},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"),
                   "reasons": clocks.get("reasons", []), "samples_in_region": 0 if clocks.get("pre_region_fallback")
                   else clocks.get("samples", 0), "pre_region_fallback": bool(clocks.get("pre_region_fallback")),
                   "source": clocks.get("source")},
        "repo_native_libs_loaded": marks.get("a_loaded", []),
        "e2e": {"value": samples / (b_ms / 1e3), "unit": "samples/s", "ms_per_step": b_ms / K,
                "h2d_bytes_per_step": B * (3 * 32 * 32 * 4 + 8), "d2h_bytes_per_step": 0,
                "api": "dpp.train(rank, world_size) with its stock DataLoader (pageable H2D, loss.item() every 100 steps)"},
        "gpu_launches": 0,
    })


# =========================================================================================
# extra workloads (BASELINE.json configs 3 and 4): ResNet-50 on ImageNet-shaped batches, GPT-2 medium
# =========================================================================================
def run_extra(args, rank, local_rank, world):
    """``--model resnet50|gpt2-medium`` for either arm.  ours: this framework (DDP + TrainStep, native kernels);
    reference: the stock stack the reference script is written against -- torchvision / HF transformers models under
    ``torch.nn.parallel.DistributedDataParallel`` over NCCL, eager, cuDNN / cuBLAS / SDPA -- none of this repo's code.
    Same JSON contract as the headline: device-timed K steps (max over ranks), clocks sampled inside the region, and an
    end-to-end figure that includes the pinned host->device copy of every step's inputs and a loss read-back."""
    import torch
    import torch.distributed as dist

    ClockSampler = load_clock_sampler()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    ours = args.impl == "ours"
    is_gpt = args.model.startswith("gpt2")
    B = args.batch or (8 if is_gpt else 32)
    dtype = args.dtype or ("bf16" if is_gpt else "tf32")
    K, W, T = args.steps, args.warmup, args.seq_len
    if world > 1 and not dist.is_initialized():
        if ours:
            import distributeddataparallel_b200 as d

            d.setup(rank, world, backend="nccl")
        else:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(0)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    pool_n = 8
    if is_gpt:
        host = [torch.randint(0, 50257, (B, T + 1), dtype=torch.int64, generator=g).pin_memory() for _ in range(pool_n)]
        h2d = B * (T + 1) * 8
    else:
        host = [(torch.randint(0, 256, (B, 224, 224, 3), dtype=torch.uint8, generator=g).pin_memory(),
                 torch.randint(0, 1000, (B,), dtype=torch.int64, generator=g).pin_memory()) for _ in range(pool_n)]
        h2d = B * 224 * 224 * 3 + B * 8
    launches_fn = lambda: 0
    if ours:
        import distributeddataparallel_b200 as d
        from distributeddataparallel_b200.ops import native

        if is_gpt:
            from distributeddataparallel_b200.models.gpt2 import build_gpt2

            model = build_gpt2(args.model, T).to(dev, torch.bfloat16 if dtype == "bf16" else torch.float32)
            crit = None
        else:
            model = d.resnet50(1000).to(dev)
            crit = d.CrossEntropyLoss()
        ddp = d.DDP(model, device_ids=[local_rank], comm=args.comm, allreduce_algo=args.algo)
        opt = d.SGD(ddp.parameters(), lr=0.01)
        step = d.TrainStep(ddp, crit, opt, cuda_graph=not args.no_graph)

        def to_device(h):
            if is_gpt:
                t = h.to(dev, non_blocking=True)
                return t[:, :-1].contiguous(), t[:, 1:].contiguous()
            return d.ops.preprocess_u8(h[0].to(dev, non_blocking=True)), h[1].to(dev, non_blocking=True)

        run = lambda x, y: step(x, y)
        comm_name = ddp.comm.name
        launches_fn = lambda: native.launch_count() + step.replays * native.launches_per_capture()
        api = "pinned batch -> H2D -> TrainStep(DDP(model), loss, SGD) -> AsyncLossReader (pinned D2H every step)"
    else:
        from torch.nn.parallel import DistributedDataParallel as StockDDP

        if is_gpt:
            from transformers import GPT2Config, GPT2LMHeadModel

            cfg = GPT2Config(n_layer=24, n_head=16, n_embd=1024, n_positions=T, vocab_size=50257, attn_pdrop=0.0,
                             embd_pdrop=0.0, resid_pdrop=0.0) if args.model == "gpt2-medium" else GPT2Config()
            cfg._attn_implementation = "sdpa"
            model = GPT2LMHeadModel(cfg).to(dev, torch.bfloat16 if dtype == "bf16" else torch.float32)
        else:
            import torchvision

            model = torchvision.models.resnet50(weights=None).to(dev).to(memory_format=torch.channels_last)
        net = StockDDP(model, device_ids=[local_rank]) if world > 1 else model
        opt = torch.optim.SGD(model.parameters(), lr=0.01)

        def to_device(h):
            if is_gpt:
                t = h.to(dev, non_blocking=True)
                return t[:, :-1].contiguous(), t[:, 1:].contiguous()
            x = h[0].to(dev, non_blocking=True).permute(0, 3, 1, 2).float().div_(255.0).sub_(0.5).div_(0.5)
            return x.contiguous(memory_format=torch.channels_last), h[1].to(dev, non_blocking=True)

        def run(x, y):
            opt.zero_grad()
            if is_gpt:
                loss = net(input_ids=x, labels=x).loss
            elif dtype == "bf16":
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    loss = torch.nn.functional.cross_entropy(net(x), y)
            else:
                loss = torch.nn.functional.cross_entropy(net(x), y)
            loss.backward()
            opt.step()
            return loss.detach()

        comm_name = "nccl (torch DDP)" if world > 1 else "none"
        api = "pinned batch -> H2D -> stock DDP(model) forward / backward / torch.optim.SGD -> pinned async D2H of the loss every step"

    pool = [to_device(h) for h in host]
    sampler = ClockSampler(gpu_index=local_rank).start()
    for i in range(4 + W):
        run(*pool[i % pool_n])
    barrier(world)
    l0 = launches_fn()
    mark = sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        loss = run(*pool[i % pool_n])
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    dev_ms = max_over_ranks(e0.elapsed_time(e1), world, dev)
    clocks = sampler.summary(mark, sampler.mark())
    launches = launches_fn() - l0
    loss_val = float(loss.item())
    # end to end: inputs come from pinned host memory every step, the loss goes back to the host every step (pinned,
    # asynchronous, consumed one step later -- the same pattern in both arms: AsyncLossReader here, a plain pinned
    # ring for the stock stack)
    if ours:
        reader = d.AsyncLossReader()
    else:
        class _Ring:
            def __init__(self):
                self.buf = torch.zeros(4, dtype=torch.float32).pin_memory()
                self.ev = [torch.cuda.Event() for _ in range(4)]
                self.q, self.n = [], 0

            def pending(self):
                return len(self.q)

            def push(self, loss):
                s = self.n % 4
                self.n += 1
                self.buf[s:s + 1].copy_(loss.detach().reshape(1).float(), non_blocking=True)
                self.ev[s].record()
                self.q.append(s)

            def pop(self):
                s = self.q.pop(0)
                self.ev[s].synchronize()
                return float(self.buf[s])

            def drain(self):
                return [self.pop() for _ in range(len(self.q))]

        reader = _Ring()
    for i in range(W):
        run(*to_device(host[i % pool_n])).item()
    barrier(world)
    t0 = time.perf_counter()
    host_loss = 0.0
    for i in range(K):
        reader.push(run(*to_device(host[i % pool_n])))
        if reader.pending() > 1:
            host_loss = reader.pop()
    for v in reader.drain():
        host_loss = v
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world, dev)
    sampler.stop()
    unit_scale = T if is_gpt else 1
    unit = "tokens/s" if is_gpt else "samples/s"
    samples = K * B * world * unit_scale
    emit(rank, {
        "metric": "tokens/sec" if is_gpt else "samples/sec", "value": samples / (dev_ms / 1e3), "unit": unit, "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": dtype, "impl": args.impl,
        "data": "synthetic " + ("token streams" if is_gpt else "ImageNet-shaped uint8 images") + ", random-init weights",
        "config": {"model": args.model, "global_batch": B * world, "per_gpu_batch": B, "seq_len": T if is_gpt else None,
                   "image": None if is_gpt else "3x224x224", "optimizer": "SGD lr=0.01", "parallelism": f"dp{world}",
                   "comm": comm_name, "cuda_graph": bool(ours and not args.no_graph),
                   "l2_policy": "model state (parameters + gradients, >= 200 MB) and activations exceed the 126 MB L2; "
                                "inputs rotate through a pool of 8 batches"},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"), "reasons": clocks.get("reasons", []),
                   "samples_in_region": 0 if clocks.get("pre_region_fallback") else clocks.get("samples", 0),
                   "pre_region_fallback": bool(clocks.get("pre_region_fallback")), "source": clocks.get("source")},
        "e2e": {"value": samples / (e2e_ms / 1e3), "unit": unit, "ms_per_step": e2e_ms / K, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "api": api},
        "gpu_launches": int(launches) if ours else 0, "final_loss": loss_val, "e2e_final_loss": host_loss,
        "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30,
        "repo_native_libs_loaded": loaded_native_libs(),
    })
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    rank, local_rank, world = dist_env(args)
    quiet_stdout()
    if args.model not in ("simplecnn", "dpp", "resnet18-cifar"):
        try:
            run_extra(args, rank, local_rank, world)
        except Exception as e:
            import traceback

            traceback.print_exc()
            emit(rank, {"impl": args.impl, "model": args.model, "unavailable": f"{type(e).__name__}: {e}"[:300]})
            if world > 1:
                sys.stdout.flush()
                os._exit(3)          # leave at once with a failure code: the launcher then stops the peers instead of
                                     # letting them wait in a collective for this rank
        return
    args.batch = args.batch or BATCH
    if args.impl == "reference":
        try:
            run_reference(args, rank, local_rank, world)
        except Exception as e:
            import traceback

            traceback.print_exc()
            emit(rank, {"impl": "reference", "unavailable": f"reference run failed: {type(e).__name__}: {e}"[:300]})
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
