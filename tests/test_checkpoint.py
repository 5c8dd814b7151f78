"""Checkpoint contract: files written here load into a stock
``torch.nn.parallel.DistributedDataParallel(SimpleCNN())`` and files written from the stock wrapper load here;
optimizer state dicts are interchangeable with ``torch.optim.SGD``."""
import os
import socket

import pytest
import torch
import torch.distributed as dist

import distributeddataparallel_b200 as d
from distributeddataparallel_b200.checkpoint import load_checkpoint, save_checkpoint

tv = pytest.importorskip("torchvision")


def _stock_simplecnn():
    class SimpleCNN(torch.nn.Module):          # /root/reference/dpp.py:11-18 with weights=None
        def __init__(self):
            super().__init__()
            self.model = tv.models.resnet18(weights=None)
            self.model.fc = torch.nn.Linear(512, 10)

        def forward(self, x):
            return self.model(x)

    return SimpleCNN()


@pytest.fixture()
def single_rank_group():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=0, world_size=1)
    d.ops.set_native(False)
    yield
    dist.destroy_process_group()


def _one_step(model, opt):
    torch.manual_seed(3)
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 10, (4,))
    opt.zero_grad()
    torch.nn.functional.cross_entropy(model(x), y).backward()
    opt.step()


def test_our_checkpoint_loads_into_stock_ddp(tmp_path, single_rank_group):
    from torch.nn.parallel import DistributedDataParallel as StockDDP

    torch.manual_seed(0)
    ours = d.DDP(d.SimpleCNN())
    opt = d.SGD(ours.parameters(), lr=0.01, momentum=0.9)
    _one_step(ours, opt)
    path = str(tmp_path / "ck.pt")
    save_checkpoint(path, ours, opt, epoch=2, step=7)
    payload = torch.load(path, weights_only=False)
    assert len(payload["model"]) == 122 and all(k.startswith("module.model.") for k in payload["model"])
    assert all(v.is_contiguous() for v in payload["model"].values())

    stock = StockDDP(_stock_simplecnn())
    missing = stock.load_state_dict(payload["model"], strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    sopt = torch.optim.SGD(stock.parameters(), lr=0.5)
    sopt.load_state_dict(payload["optimizer"])
    assert sopt.param_groups[0]["lr"] == 0.01 and sopt.param_groups[0]["momentum"] == 0.9
    for a, b in zip(ours.state_dict().values(), stock.state_dict().values()):
        assert torch.equal(a.contiguous(), b)
    # momentum buffers came across: one more identical step keeps the replicas identical
    _one_step(ours, opt)
    _one_step(stock, sopt)
    for (k, a), b in zip(ours.state_dict().items(), stock.state_dict().values()):
        assert torch.allclose(a.float(), b.float(), rtol=1e-4, atol=1e-5), k


def test_stock_checkpoint_loads_here(tmp_path, single_rank_group):
    from torch.nn.parallel import DistributedDataParallel as StockDDP

    torch.manual_seed(1)
    stock = StockDDP(_stock_simplecnn())
    sopt = torch.optim.SGD(stock.parameters(), lr=0.02)
    _one_step(stock, sopt)
    path = str(tmp_path / "stock.pt")
    torch.save({"model": stock.state_dict(), "optimizer": sopt.state_dict(), "epoch": 1, "step": 3}, path)

    ours = d.DDP(d.SimpleCNN())
    opt = d.SGD(ours.parameters(), lr=0.9)
    info = load_checkpoint(path, ours, opt)
    assert info["epoch"] == 1 and info["step"] == 3 and opt.param_groups[0]["lr"] == 0.02
    for a, b in zip(ours.state_dict().values(), stock.state_dict().values()):
        assert torch.equal(a.contiguous(), b)

    # an unwrapped model accepts the same file (the "module." prefix is stripped) and vice versa
    bare = d.SimpleCNN()
    load_checkpoint(path, bare)
    assert torch.equal(bare.model.fc.weight, stock.module.model.fc.weight)
    bare_path = str(tmp_path / "bare.pt")
    torch.save(bare.state_dict(), bare_path)                    # a plain state_dict, no wrapper keys
    again = d.DDP(d.SimpleCNN())
    load_checkpoint(bare_path, again)
    assert torch.equal(again.module.model.fc.weight, stock.module.model.fc.weight)
