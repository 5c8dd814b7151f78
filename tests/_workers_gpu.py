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
            for algo in ("oneshot", "twoshot"):
                if algo == "oneshot" and numel * 4 > (4 << 20):
                    continue
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

    # one eager step on both wrappers from identical state: averaged grads must agree with NCCL
    x, y = batch()
    d.CrossEntropyLoss()(model(x), y).backward()
    d.CrossEntropyLoss()(ref_model(x), y).backward()
    ga = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    gb = torch.cat([p.grad.reshape(-1) for p in ref_model.parameters()])
    rec = {"grad_err_vs_nccl": float((ga - gb).norm() / gb.norm())}
    opt.zero_grad()
    for _ in range(8):
        step(*batch())
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
            def unfused():
                native.gemm_tc(dy.t().contiguous() if False else dy, x, 2, out=None) if False else None
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
    comm.check()
    rec["tc_error"] = native.tc_error_word(dev)
    torch.save(rec, os.path.join(out_dir, f"fused{rank}.pt"))
    d.cleanup()


def nvls_worker(rank, ws, port, out_dir):
    """NVLS (multimem) all-reduce through the symm_mem-provided arena, where the box supports multicast."""
    d = _init(rank, ws, port)
    from distributeddataparallel_b200.parallel.p2p import P2PComm

    dev = torch.device("cuda", rank)
    rec = {"ok": True, "fails": [], "multicast": False, "times": {}}
    try:
        comm = P2PComm(None, dev, algo="nvls")
        comm.reserve(128 << 20)
    except Exception as e:                      # symm_mem allocator unavailable on this box
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
            for algo in ("nvls", "twoshot"):        # both kernels must work on the symm_mem arena
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
