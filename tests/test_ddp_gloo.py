"""Distributed tier on CPU (gloo, world_size 2, mp.spawn): BASELINE.json config #1."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

import _workers as W


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("channels_last", [False, True])
def test_parity_with_stock_ddp(tmp_path, channels_last):
    ws, steps = 2, 3
    mp.spawn(W.ddp_parity_worker, args=(ws, _port(), str(tmp_path), steps, channels_last), nprocs=ws, join=True)
    recs = [torch.load(tmp_path / f"rank{r}.pt") for r in range(ws)]
    for r in recs:
        assert r["ctor_equal"], "constructor broadcast must equalise parameters"
        assert all(r["grads_identical"]), "averaged gradients must be bitwise identical on every rank"
        if channels_last:
            # NHWC and NCHW CPU conv kernels round differently and batch-8 BatchNorm on 1x1 maps
            # amplifies that chaotically after a few updates: only the first step is comparable.
            assert r["grad_err"][0] < 1e-3 and abs(r["loss_a"][0] - r["loss_b"][0]) < 1e-4
        else:
            assert max(r["grad_err"]) < 2e-3
            for a, b in zip(r["loss_a"], r["loss_b"]):
                assert abs(a - b) < 1e-3 * max(1.0, abs(b))
            assert r["param_err"] < 1e-4
        # iteration 0: one bucket; after the one-time rebuild: three
        assert r["buckets"] == [1, 3, 3]
        log = r["log"]
        assert log["has_rebuilt_buckets"] and log["num_parameter_tensors"] == 62
        assert log["total_parameter_size_bytes"] == 44_726_568
        assert log["initial_bucket_sizes"] == [44_726_568]
        assert log["rebuilt_bucket_sizes"] == [9_461_800, 26_494_976, 8_769_792]
        assert len(r["keys"]) == 122 and all(k.startswith("module.model.") for k in r["keys"])
        assert r["keys"][0] == "module.model.conv1.weight" and r["keys"][-1] == "module.model.fc.bias"


def test_streamed_last_bucket_keeps_parity_with_stock_ddp(tmp_path):
    """tail_splits: the last bucket is reduced in three pieces (5 collectives per step instead of 3); gradients,
    losses and parameters still match stock DDP and the reported layout stays the stock one."""
    ws, steps = 2, 3
    mp.spawn(W.ddp_parity_worker, args=(ws, _port(), str(tmp_path), steps, False, [1 << 20, 64 << 10]), nprocs=ws, join=True)
    for rank in range(ws):
        r = torch.load(tmp_path / f"rank{rank}.pt")
        assert all(r["grads_identical"]) and max(r["grad_err"]) < 2e-3 and r["param_err"] < 1e-4
        assert r["buckets"] == [1, 5, 5]
        log = r["log"]
        assert log["rebuilt_bucket_sizes"] == [9_461_800, 26_494_976, 8_769_792]
        assert sum(log["launch_unit_sizes"]) == 44_726_568 and len(log["launch_unit_sizes"]) == 5
        assert log["launch_unit_sizes"][-1] <= (64 << 10) and log["num_buckets_reduced"] == 1 + 3 * 2


def test_ddp_features(tmp_path):
    ws = 2
    mp.spawn(W.features_worker, args=(ws, _port(), str(tmp_path)), nprocs=ws, join=True)
    for r in range(ws):
        rec = torch.load(tmp_path / f"rank{r}.pt")
        assert rec["nosync_launches"] >= 1 and rec["nosync_err"] < 1e-5
        assert rec["hook_identical"] and "compress" in rec["hook_name"]
        assert rec["unused_grad_zero"] and rec["unused_identical"]
        assert "find_unused_parameters" in rec["unfinished_error"]
        assert "same model across all ranks" in rec["mismatch_error"]


def test_train_and_resume(tmp_path):
    ws = 2
    ckpt = str(tmp_path / "ck.pt")
    mp.spawn(W.train_worker, args=(ws, _port(), str(tmp_path), ckpt, None), nprocs=ws, join=True)
    s0 = torch.load(tmp_path / "train0.pt")
    assert s0["steps"] == 6 and len(s0["loss_history"]) == 3
    assert os.path.exists(ckpt)
    payload = torch.load(ckpt, weights_only=False)
    assert len(payload["model"]) == 122 and payload["step"] == 6
    assert payload["optimizer"]["param_groups"][0]["lr"] == 0.01
    assert payload["optimizer"]["param_groups"][0]["params"] == list(range(62))
    mp.spawn(W.train_worker, args=(ws, _port(), str(tmp_path), ckpt, ckpt), nprocs=ws, join=True)
    s1 = torch.load(tmp_path / "train0.pt")
    assert s1["steps"] >= 7


def test_train_language_model_flag_path(tmp_path):
    """--model gpt2-tiny: train() feeds token batches (vocabulary of the chosen model) to the fused-loss GPT-2."""
    ws = 2
    mp.spawn(W.lm_train_worker, args=(ws, _port(), str(tmp_path)), nprocs=ws, join=True)
    s0 = torch.load(tmp_path / "lm0.pt")
    assert s0["steps"] == 4 and len(s0["loss_history"]) == 4
    assert all(l == l and 0 < l < 20 for l in s0["loss_history"])
    assert s0["ddp"]["world_size"] == 2 and s0["ddp"]["num_buckets_reduced"] >= 4


def test_train_rejects_flag_combinations_it_cannot_honour():
    import distributeddataparallel_b200 as d

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(_port())
    with pytest.raises(ValueError, match="bf16"):
        d.train(0, 1, 1, d.TrainConfig(backend="gloo", device="cpu", dtype="bf16", max_steps=1))
    d.cleanup()
    with pytest.raises(ValueError, match="seq-len"):
        d.train(0, 1, 1, d.TrainConfig(backend="gloo", device="cpu", model="gpt2-tiny", seq_len=4096, max_steps=1))
    d.cleanup()


@pytest.mark.parametrize("divide_by_world", [True, False])
def test_join_uneven_inputs_matches_stock(tmp_path, divide_by_world):
    """DDP.join(): rank 0 runs 2 iterations, rank 1 runs 5; nobody hangs, replicas end identical and equal
    to what stock DDP's join() produces on the same data -- with gradients divided by the world size or by the
    number of ranks that still have data."""
    ws = 2
    mp.spawn(W.join_worker, args=(ws, _port(), str(tmp_path), divide_by_world), nprocs=ws, join=True)
    for r in range(ws):
        rec = torch.load(tmp_path / f"join{r}.pt")
        assert rec["identical_across_ranks"]
        assert rec["err_vs_stock"] < 1e-5, rec
        assert rec["control"] >= rec["iters"] + 1
        assert rec["divisor_after"] == ws


def test_join_throw_on_early_termination(tmp_path):
    ws = 2
    mp.spawn(W.join_throw_worker, args=(ws, _port(), str(tmp_path)), nprocs=ws, join=True)
    recs = [torch.load(tmp_path / f"throw{r}.pt") for r in range(ws)]
    assert all("exhausted inputs" in r["msg"] for r in recs), recs
    assert recs[0]["done"] == 2 and recs[1]["done"] == 2        # rank 1 stops in the iteration rank 0 misses


def test_params_and_buffers_to_ignore(tmp_path):
    ws = 2
    mp.spawn(W.ignore_worker, args=(ws, _port(), str(tmp_path)), nprocs=ws, join=True)
    for r in range(ws):
        rec = torch.load(tmp_path / f"ign{r}.pt")
        assert rec["kept_local"] and rec["ignored_grad_differs"] and rec["other_grad_equal"], rec
        assert rec["ignored_buffer_differs"] and rec["will_sync"]
        assert "unused.weight" not in rec["names"] and "unused.bias" in rec["names"]


def test_fault_injection_fail_fast():
    ws = 2
    with pytest.raises(Exception) as ei:
        mp.spawn(W.failing_worker, args=(ws, _port()), nprocs=ws, join=True)
    assert "injected failure on rank 1" in str(ei.value)
