"""Collective + integration tiers on >= 2 real GPUs (run with `gpurun --gpus N`)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

import _workers_gpu as W

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _n():
    return min(torch.cuda.device_count(), int(os.environ.get("B200DDP_TEST_GPUS", 8)))


def test_p2p_collectives_match_nccl(tmp_path):
    n = _n()
    mp.spawn(W.collectives_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    for r in range(n):
        rec = torch.load(tmp_path / f"coll{r}.pt")
        assert rec["ok"], rec


def test_ddp_p2p_end_to_end(tmp_path):
    n = _n()
    mp.spawn(W.ddp_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    recs = [torch.load(tmp_path / f"ddp{r}.pt") for r in range(n)]
    for rec in recs:
        assert rec["comm"] == "p2p"
        assert rec["params_identical_across_ranks"] and rec["buffers_identical_after_sync"]
        # identical local gradients reduced by the product kernels (real launch units) and by NCCL: rounding only
        assert rec["grad_err_vs_nccl"] < 2e-6, rec["grad_err_vs_nccl"]
        assert rec["graph_replays"] >= 3 and rec["error_word"] == 0
        assert rec["rebuilt"] == [9_461_800, 26_494_976, 8_769_792]
        # the last bucket is streamed: its tail unit is the stem's 37 632-byte weight gradient (+ bn1), pushed
        assert len(rec["launch_units"]) > 3 and sum(rec["launch_units"]) == 44_726_568 and rec["launch_units"][-1] <= 65_536
        assert sum(rec["algo_counts"].values()) >= len(rec["launch_units"])
        assert rec["graphs"] == 2            # steady-state batch + ragged tail
    print("ddp:", {k: recs[0][k] for k in ("provider", "multicast", "algo_counts", "launch_units")})


def test_fused_wgrad_allreduce(tmp_path):
    n = min(_n(), 4)
    mp.spawn(W.fused_wgrad_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    for r in range(n):
        rec = torch.load(tmp_path / f"fused{r}.pt")
        assert rec["ok"] and rec["tc_error"] == 0, rec
    rec0 = torch.load(tmp_path / "fused0.pt")
    print("fused vs unfused ms:", rec0["times"])
    print("NVLS-fused GEMM vs GEMM + all-reduce (ms):", rec0["nvls"])
    assert not rec0["nvls"]["fails"], rec0["nvls"]


def test_builtin_comm_hooks_are_fused_on_the_p2p_plane(tmp_path):
    n = _n()
    mp.spawn(W.hooks_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    for r in range(n):
        rec = torch.load(tmp_path / f"hooks{r}.pt")
        for name, v in rec.items():
            assert v["fused"] > 0 and v["same"] and v["err"] < v["tol"], (name, v)
        assert rec["bf16"]["algos"].get("wire16", 0) > 0, rec


def test_dead_rank_makes_survivors_raise(tmp_path):
    """SIGKILL-style death of one rank mid-training: every survivor raises RuntimeError within a bounded time and its
    GPU still answers (time-bounded waits + NaN poison + host-polled error word)."""
    import multiprocessing

    n = _n()
    port = _port()
    ctx = multiprocessing.get_context("spawn")
    procs = [ctx.Process(target=W.dead_rank_worker, args=(r, n, port, str(tmp_path))) for r in range(n)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.kill()
    assert not alive, "survivors hung"
    assert procs[-1].exitcode == 17
    for r in range(n - 1):
        rec = torch.load(tmp_path / f"dead{r}.pt")
        assert rec["raised"] and rec["seconds"] < 60 and rec["gpu_alive"], rec
        assert "timed out" in rec["message"]


def test_nvls_allreduce_where_multicast_exists(tmp_path):
    n = _n()
    mp.spawn(W.nvls_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    rec = torch.load(tmp_path / "nvls0.pt")
    print("nvls:", {k: v for k, v in rec.items() if k != "fails"})
    if "unavailable" in rec:
        pytest.skip("symmetric arena unavailable: " + rec["unavailable"])
    if not rec["multicast"]:
        pytest.skip("this box exposes no NVLink multicast")
    for r in range(n):
        assert torch.load(tmp_path / f"nvls{r}.pt")["ok"], torch.load(tmp_path / f"nvls{r}.pt")
