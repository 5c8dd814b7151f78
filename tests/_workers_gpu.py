"""mp.spawn workers for the multi-GPU tests."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist


def _init(rank, ws, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import distributeddataparallel_b200 as d

    torch.cuda.set_device(rank)
    d.setup(rank, ws, backend="nccl")
    return d


def collectives_worker(rank, ws, port, out_dir):
    d = _init(rank, ws, port)
    from distributeddataparallel_b200.parallel.p2p import P2PComm

    dev = torch.device("cuda", rank)
    comm = P2PComm(None, dev)
    comm.reserve(64 << 20)
    rec = {"ok": True, "fails": []}
    g = torch.Generator(device="cpu").manual_seed(rank)
    for dtype in (torch.float32, torch.bfloat16):
        for numel in (256, 65_536, 1 << 20, 5_000_000):
            for algo in ("oneshot", "twoshot", "push", "auto", "nvls"):
                if algo == "oneshot" and numel * 4 > (4 << 20):
                    continue
                if algo == "push" and numel * 4 > comm.push_max:
                    continue
                if algo == "nvls" and not comm.arena.multicast:
                    continue
                if numel == 65_536 and algo == "twoshot" and dtype == torch.float32:
                    comm.arena.seed_epochs(2 ** 32 - 6)          # cross the 2^32 epoch boundary on real peers
                    dist.barrier()
                comm.algo = algo
                flat = comm.allocate(numel, dtype, dev)
                src = torch.randn(numel, generator=g).to(dtype).to(dev)
                flat.copy_(src)
                ref = src.float().clone()
                dist.all_reduce(ref)
                ref /= ws
                comm.launch(0, flat, True).wait()
                torch.cuda.synchronize()
                gathered = [torch.empty_like(flat) for _ in range(ws)]
                dist.all_gather(gathered, flat)
                same = all(torch.equal(gathered[0], t) for t in gathered)
                tol = 1e-5 if dtype == torch.float32 else 2e-2
                close = torch.allclose(flat.float(), ref, rtol=tol, atol=tol)
                if not (same and close):
                    rec["ok"] = False
                    rec["fails"].append((str(dtype), numel, algo, same, close))
    # broadcast of a tensor outside the arena (staged) and the barrier
    t = torch.full((1_000_003,), float(rank), device=dev)
    comm.broadcast_(t, src=ws - 1)
    comm.barrier()
    torch.cuda.synchronize()
    if not bool(torch.all(t == float(ws - 1))):
        rec["ok"] = False
        rec["fails"].append("broadcast")
    comm.check()
    rec["provider"] = comm.arena.provider
    rec["multicast"] = bool(comm.arena.multicast)
    rec["multicast_error"] = getattr(comm.arena, "multicast_error", None)
    rec["vmm_error"] = getattr(comm.arena, "vmm_error", None)
    torch.save(rec, os.path.join(out_dir, f"coll{rank}.pt"))
    d.cleanup()


def ddp_worker(rank, ws, port, out_dir):
    d = _init(rank, ws, port)
    dev = torch.device("cuda", rank)
    torch.manual_seed(100 + rank)                      # constructor broadcast must equalise
    model = d.DDP(d.SimpleCNN().to(dev), device_ids=[rank], comm="p2p")
    ref_model = d.DDP(d.SimpleCNN().to(dev), device_ids=[rank], comm="c10d")
    ref_model.load_state_dict(model.state_dict())
    opt = d.SGD(model.parameters(), lr=0.01)
    step = d.TrainStep(model, d.CrossEntropyLoss(), opt, cuda_graph=True)
    g = torch.Generator(device="cpu").manual_seed(7 + rank)

    def batch():
        u = torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
        return d.ops.preprocess_u8(u), torch.randint(0, 10, (32,), generator=g).to(dev)

    # two eager steps (the second one runs on the rebuilt, tail-split bucket layout)
    for _ in range(2):
        x, y = batch()
        opt.zero_grad()
        d.CrossEntropyLoss()(model(x), y).backward()
    # local gradients of one more batch, NOT reduced (no_sync): reduce exactly these with the product kernels through
    # the real launch units, and with NCCL, and compare -- identical inputs, so the tolerance is rounding only
    opt.zero_grad()
    x, y = batch()
    with model.no_sync():
        d.CrossEntropyLoss()(model(x), y).backward()
    torch.cuda.synchronize()
    units = [model.reducer.bucket_flat(i) for i in range(model.reducer.num_buckets())]
    expect = []
    for flat in units:
        e = flat.detach().clone()
        dist.all_reduce(e)
        expect.append(e / ws)
    for i, flat in enumerate(units):
        model.comm.launch(i, flat, i + 1 == len(units)).wait()
    torch.cuda.synchronize()
    rec = {"grad_err_vs_nccl": max(float((f - e).abs().max() / e.abs().max().clamp_min(1e-20))
                                   for f, e in zip(units, expect)),
           "launch_units": model._get_ddp_logging_data()["launch_unit_sizes"],
           "algo_counts": dict(model.comm.algo_counts), "provider": model.comm.arena.provider,
           "multicast": bool(model.comm.arena.multicast)}
    del ref_model
    opt.zero_grad()
    for _ in range(8):
        step(*batch())
    # the ragged tail: first sight eager, second sight captured as a second graph
    def tail():
        u = torch.randint(0, 256, (16, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
        return d.ops.preprocess_u8(u), torch.randint(0, 10, (16,), generator=g).to(dev)
    step(*tail())
    step(*batch())
    step(*tail())
    step(*batch())
    rec["graphs"] = len(step._graphs)
    torch.cuda.synchronize()
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(ws)]
    dist.all_gather(gathered, flat)
    rec["params_identical_across_ranks"] = all(torch.equal(gathered[0], t) for t in gathered)
    model._sync_buffers()
    torch.cuda.synchronize()
    bflat = torch.cat([b.detach().float().reshape(-1) for b in model.buffers()])
    gb2 = [torch.empty_like(bflat) for _ in range(ws)]
    dist.all_gather(gb2, bflat)
    rec["buffers_identical_after_sync"] = all(torch.equal(gb2[0], t) for t in gb2)
    rec["graph_replays"] = step.replays
    rec["comm"] = model.comm.name
    rec["rebuilt"] = model._get_ddp_logging_data()["rebuilt_bucket_sizes"]
    model.comm.check()
    rec["error_word"] = model.comm.arena.error_word()
    torch.save(rec, os.path.join(out_dir, f"ddp{rank}.pt"))
    del step
    d.cleanup()


def fused_wgrad_worker(rank, ws, port, out_dir):
    """GEMM + all-reduce in one kernel vs GEMM followed by the all-reduce kernel."""
    d = _init(rank, ws, port)
    from distributeddataparallel_b200.ops import native
    from distributeddataparallel_b200.parallel.p2p import FusedGradReduce, P2PComm

    dev = torch.device("cuda", rank)
    comm = P2PComm(None, dev)
    comm.reserve(512 << 20)
    fused = FusedGradReduce(comm)
    rec = {"ok": True, "fails": [], "times": {}}
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    for dtype in (torch.bfloat16, torch.float32):
        for (tok, n_out, k_in) in [(1024, 256, 128), (4096, 1024, 1024), (8192, 4096, 1024)]:
            dy = torch.randn(tok, n_out, generator=g).to(dtype).to(dev)
            x = torch.randn(tok, k_in, generator=g).to(dtype).to(dev)
            out = comm.allocate(n_out * k_in, dtype, dev).view(n_out, k_in)
            ref = (dy.double().t() @ x.double())
            dist.all_reduce(ref)
            ref /= ws
            for rep in range(3):                         # repeated launches reuse staging / flags
                out.zero_()
                torch.cuda.synchronize(); dist.barrier()
                fused.wgrad(dy, x, out)
            torch.cuda.synchronize()
            gathered = [torch.empty_like(out) for _ in range(ws)]
            dist.all_gather(gathered, out)
            same = all(torch.equal(gathered[0], t) for t in gathered)
            err = float((out.double() - ref).norm() / ref.norm())
            tol = 2e-2 if dtype == torch.bfloat16 else 3e-3
            if not (same and err < tol):
                rec["ok"] = False
                rec["fails"].append((str(dtype), tok, n_out, k_in, same, err))
            # timing: fused vs unfused (GEMM into the arena slot, then the two-shot all-reduce)
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            from distributeddataparallel_b200.ops import native_tf
            tmp = comm.allocate(n_out * k_in, torch.float32, dev).view(n_out, k_in) if dtype == torch.float32 else \
                comm.allocate(n_out * k_in, dtype, dev).view(n_out, k_in)
            torch.cuda.synchronize(); dist.barrier()
            e0.record()
            for _ in range(5):
                fused.wgrad(dy, x, out)
            e1.record()
            for _ in range(5):
                native_tf.gemm_ex(dy, x, tmp, 2, n_out, k_in, tok, n_out, k_in, k_in)
                comm.all_reduce_avg_(tmp.view(-1))
            e2.record()
            torch.cuda.synchronize()
            rec["times"][(str(dtype), tok, n_out, k_in)] = (e0.elapsed_time(e1) / 5, e1.elapsed_time(e2) / 5)
    # ---- NVLS form: persistent 2-CTA GEMM + in-switch reduction in one kernel (bf16, multicast arena) ----
    rec["nvls"] = {"available": bool(comm.arena.multicast), "times": {}, "fails": []}
    if comm.arena.multicast:
        from distributeddataparallel_b200.ops import native_tf

        for (tok, n_out, k_in) in [(4096, 1024, 1024), (8192, 4096, 1024), (8192, 1024, 4096), (8192, 3072, 1024)]:
            dy = torch.randn(tok, n_out, generator=g).to(torch.bfloat16).to(dev)
            x = torch.randn(tok, k_in, generator=g).to(torch.bfloat16).to(dev)
            out = comm.allocate(n_out * k_in, torch.bfloat16, dev).view(n_out, k_in)
            tmp = comm.allocate(n_out * k_in, torch.bfloat16, dev).view(n_out, k_in)
            # reference with the same rounding points: per-rank product rounded to bf16, summed in fp32, scaled, rounded
            part = (dy.float().t() @ x.float()).to(torch.bfloat16).float()
            dist.all_reduce(part)
            ref = (part / ws).to(torch.bfloat16)
            ok_shape = True
            for rep in range(3):
                out.fill_(float("nan"))
                torch.cuda.synchronize(); dist.barrier()
                ok_shape = fused.wgrad_nvls(dy, x, out)
                if not ok_shape:
                    break
            torch.cuda.synchronize()
            if not ok_shape:
                rec["nvls"]["fails"].append(("not eligible", tok, n_out, k_in))
                continue
            gathered = [torch.empty_like(out) for _ in range(ws)]
            dist.all_gather(gathered, out)
            same = all(torch.equal(gathered[0], t) for t in gathered)
            err = float((out.float() - ref.float()).norm() / ref.float().norm())
            if not (same and err < 1e-2 and bool(torch.isfinite(out).all())):
                rec["ok"] = False
                rec["nvls"]["fails"].append((tok, n_out, k_in, same, err))
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            torch.cuda.synchronize(); dist.barrier()
            e0.record()
            for _ in range(10):
                fused.wgrad_nvls(dy, x, out)
            e1.record()
            for _ in range(10):
                native_tf.gemm_ex(dy, x, tmp, 2, n_out, k_in, tok, n_out, k_in, k_in)
                comm.all_reduce_avg_(tmp.view(-1))
            e2.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / 10, e1.elapsed_time(e2) / 10], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            rec["nvls"]["times"][(tok, n_out, k_in)] = tuple(round(float(v), 4) for v in t)
    comm.check()
    rec["tc_error"] = native.tc_error_word(dev)
    torch.save(rec, os.path.join(out_dir, f"fused{rank}.pt"))
    d.cleanup()


def hooks_worker(rank, ws, port, out_dir):
    """Built-in comm hooks on the P2P data plane run as fused kernels (16-bit wire, fp32 accumulate)."""
    d = _init(rank, ws, port)
    from distributeddataparallel_b200.parallel import comm as hooks

    dev = torch.device("cuda", rank)
    rec = {}
    for name, hook, tol in (("bf16", hooks.bf16_compress_hook, 3e-2), ("fp16", hooks.fp16_compress_hook, 5e-3),
                            ("allreduce", hooks.allreduce_hook, 1e-5)):
        torch.manual_seed(5)
        model = d.DDP(d.SimpleCNN().to(dev), device_ids=[rank], comm="p2p")
        model.register_comm_hook(None, hook)
        g = torch.Generator(device="cpu").manual_seed(7 + rank)
        x = d.ops.preprocess_u8(torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev))
        y = torch.randint(0, 10, (32,), generator=g).to(dev)
        with model.no_sync():
            d.CrossEntropyLoss()(model(x), y).backward()
        torch.cuda.synchronize()
        units = [model.reducer.bucket_flat(i) for i in range(model.reducer.num_buckets())]
        expect = []
        for flat in units:
            e = flat.detach().clone()
            dist.all_reduce(e)
            expect.append(e / ws)
        for i, flat in enumerate(units):
            model.comm.launch(i, flat, i + 1 == len(units)).wait()
        torch.cuda.synchronize()
        err = max(float((f - e).abs().max() / e.abs().max().clamp_min(1e-20)) for f, e in zip(units, expect))
        same = True
        for flat in units:
            gathered = [torch.empty_like(flat) for _ in range(ws)]
            dist.all_gather(gathered, flat)
            same = same and all(torch.equal(gathered[0], t) for t in gathered)
        rec[name] = {"err": err, "tol": tol, "same": same, "fused": getattr(model.comm, "fused_launches", 0),
                     "algos": dict(model.comm.inner.algo_counts)}
        model.check_health(sync=True)
        del model
    torch.save(rec, os.path.join(out_dir, f"hooks{rank}.pt"))
    d.cleanup()


def dead_rank_worker(rank, ws, port, out_dir):
    """The last rank dies mid-training; the survivors must raise within a bounded time (no hung GPU)."""
    import time

    os.environ["B200DDP_COMM_TIMEOUT_S"] = "3"
    d = _init(rank, ws, port)
    dev = torch.device("cuda", rank)
    torch.manual_seed(1)
    model = d.DDP(d.SimpleCNN().to(dev), device_ids=[rank], comm="p2p")
    opt = d.SGD(model.parameters(), lr=0.01)
    step = d.TrainStep(model, d.CrossEntropyLoss(), opt, cuda_graph=True)
    g = torch.Generator(device="cpu").manual_seed(rank)
    x = d.ops.preprocess_u8(torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev))
    y = torch.randint(0, 10, (32,), generator=g).to(dev)
    for _ in range(8):
        step(x, y)
    torch.cuda.synchronize()
    dist.barrier()
    if rank == ws - 1:
        os._exit(17)                                   # SIGKILL-like: no clean-up, no goodbye
    t0 = time.time()
    rec = {"raised": False}
    try:
        for _ in range(200):
            step(x, y).item()
            time.sleep(0.01)
    except RuntimeError as e:
        rec.update(raised=True, seconds=time.time() - t0, message=str(e)[:200])
    torch.cuda.synchronize()
    rec["gpu_alive"] = bool(torch.ones(4, device=dev).sum().item() == 4)
    torch.save(rec, os.path.join(out_dir, f"dead{rank}.pt"))
    os._exit(0)                                        # the process group has a dead member: skip the handshake


def nvls_worker(rank, ws, port, out_dir):
    """NVLS (multimem) all-reduce through this framework's own multicast-bound arena (csrc/runtime/vmm_arena.cpp)."""
    d = _init(rank, ws, port)
    from distributeddataparallel_b200.parallel.p2p import P2PComm

    dev = torch.device("cuda", rank)
    rec = {"ok": True, "fails": [], "multicast": False, "times": {}}
    try:
        comm = P2PComm(None, dev, algo="nvls")
        comm.reserve(128 << 20)
    except Exception as e:                      # neither the VMM nor the IPC arena could be set up
        rec["unavailable"] = repr(e)[:300]
        torch.save(rec, os.path.join(out_dir, f"nvls{rank}.pt"))
        d.cleanup()
        return
    rec["multicast"] = bool(comm.arena.multicast)
    rec["provider"] = comm.arena.provider
    g = torch.Generator(device="cpu").manual_seed(rank)
    for dtype in (torch.float32, torch.bfloat16):
        for numel in (4096, 1 << 20, 8_000_000):
            flat = comm.allocate(numel, dtype, dev)
            src = torch.randn(numel, generator=g).to(dtype).to(dev)
            ref = src.float().clone()
            dist.all_reduce(ref)
            ref /= ws
            for algo in ("nvls", "twoshot"):        # both kernels must work on the multicast-bound arena
                comm.algo = algo
                flat.copy_(src)
                comm.launch(0, flat, True).wait()
                torch.cuda.synchronize()
                gathered = [torch.empty_like(flat) for _ in range(ws)]
                dist.all_gather(gathered, flat)
                same = all(torch.equal(gathered[0], t) for t in gathered)
                tol = 1e-5 if dtype == torch.float32 else 2e-2
                close = torch.allclose(flat.float(), ref, rtol=tol, atol=tol)
                if not (same and close):
                    rec["ok"] = False
                    rec["fails"].append((str(dtype), numel, algo, same, close))
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize(); dist.barrier()
                e0.record()
                for _ in range(10):
                    comm.launch(0, flat, True).wait()
                e1.record()
                torch.cuda.synchronize()
                rec["times"][(str(dtype), numel, algo)] = e0.elapsed_time(e1) / 10 * 1e3
    comm.check()
    torch.save(rec, os.path.join(out_dir, f"nvls{rank}.pt"))
    d.cleanup()
