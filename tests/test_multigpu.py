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
        assert rec["grad_err_vs_nccl"] < 0.1, rec["grad_err_vs_nccl"]   # two separate forward passes: BN float-atomic noise, amplified
        assert rec["graph_replays"] >= 3 and rec["error_word"] == 0
        assert rec["rebuilt"] == [9_461_800, 26_494_976, 8_769_792]


def test_fused_wgrad_allreduce(tmp_path):
    n = min(_n(), 4)
    mp.spawn(W.fused_wgrad_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    for r in range(n):
        rec = torch.load(tmp_path / f"fused{r}.pt")
        assert rec["ok"] and rec["tc_error"] == 0, rec
    print("fused vs unfused ms:", torch.load(tmp_path / "fused0.pt")["times"])


def test_nvls_allreduce_where_multicast_exists(tmp_path):
    n = _n()
    mp.spawn(W.nvls_worker, args=(n, _port(), str(tmp_path)), nprocs=n, join=True)
    rec = torch.load(tmp_path / "nvls0.pt")
    print("nvls:", {k: v for k, v in rec.items() if k != "fails"})
    if "unavailable" in rec:
        pytest.skip("symmetric-memory allocator unavailable: " + rec["unavailable"])
    for r in range(n):
        assert torch.load(tmp_path / f"nvls{r}.pt")["ok"], torch.load(tmp_path / f"nvls{r}.pt")
