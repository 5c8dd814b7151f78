"""Module-tree / state_dict parity with the reference model."""
import pytest
import torch

import distributeddataparallel_b200 as d

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


def test_state_dict_contract():
    ours, stock = d.SimpleCNN(), _stock_simplecnn()
    a, b = ours.state_dict(), stock.state_dict()
    assert list(a.keys()) == list(b.keys()) and len(a) == 122
    for k in a:
        assert a[k].shape == b[k].shape and a[k].dtype == b[k].dtype, k
    assert sum(v.dtype == torch.int64 for v in a.values()) == 20
    assert sum(p.numel() for p in ours.parameters()) == 11_181_642
    assert [n for n, _ in ours.named_parameters()] == [n for n, _ in stock.named_parameters()]


def test_seeded_init_matches_stock():
    torch.manual_seed(0)
    ours = d.SimpleCNN()
    torch.manual_seed(0)
    stock = _stock_simplecnn()
    for (k, a), (_, b) in zip(ours.state_dict().items(), stock.state_dict().items()):
        assert torch.equal(a, b), k


@pytest.mark.parametrize("batch", [32, 10])
def test_forward_backward_matches_stock_cpu(batch):
    torch.manual_seed(0)
    ours, stock = d.SimpleCNN(), _stock_simplecnn()
    stock.load_state_dict(ours.state_dict())
    x = torch.randn(batch, 3, 32, 32)
    y = torch.randint(0, 10, (batch,))
    la = d.CrossEntropyLoss()(ours(x), y)
    lb = torch.nn.functional.cross_entropy(stock(x), y)
    assert torch.allclose(la, lb, rtol=1e-5, atol=1e-6)
    la.backward()
    lb.backward()
    for (n, p), (_, q) in zip(ours.named_parameters(), stock.named_parameters()):
        err = (p.grad - q.grad).norm() / q.grad.norm().clamp_min(1e-12)
        assert err < 2e-3, (n, float(err))
    for (n, p), (_, q) in zip(ours.named_buffers(), stock.named_buffers()):
        assert torch.allclose(p.float(), q.float(), rtol=1e-5, atol=1e-6), n


def test_resnet50_keys():
    a = d.resnet50().state_dict()
    b = tv.models.resnet50(weights=None).state_dict()
    assert list(a.keys()) == list(b.keys())
    assert all(a[k].shape == b[k].shape for k in a)


def test_ddp_stats_summary_formatting():
    from distributeddataparallel_b200.train import format_ddp_stats

    log = {"world_size": 8, "backend_name": "nccl", "comm_backend": "p2p", "comm_hook": "builtin",
           "num_parameter_tensors": 62, "total_parameter_size_bytes": 44726568, "initial_bucket_sizes": [44726568],
           "rebuilt_bucket_sizes": [9461800, 26494976, 8769792], "iteration": 10, "num_buckets_reduced": 28,
           "collectives": {"broadcast": 3}, "avg_forward_compute_time_ms": 0.4, "avg_backward_compute_time_ms": 0.6,
           "avg_backward_comm_time_ms": 0.5, "avg_backward_compute_comm_overlap_time_ms": 0.45,
           "comm_stats": {"avg_bucket_allreduce_us": 57.0, "timed_buckets": 27, "bus_bandwidth_GBs": 455.0}}
    text = format_ddp_stats(log, 2560, 0.01)
    assert "256000 samples/s" in text and "data plane p2p" in text
    assert "rebuilt [9461800, 26494976, 8769792]" in text and "iterations 10" in text
    assert "0.050 ms not overlapped" in text and "bus bandwidth 455 GB/s" in text


def test_shipped_native_libraries_match_their_sources():
    """The in-tree binaries travel to the GPU box with the snapshot; a library older than its sources would be rebuilt
    there by every rank at once.  (Skipped before the first build.)"""
    from distributeddataparallel_b200 import _build

    if not _build.KERNEL_LIB.exists():
        pytest.skip("kernel library not built yet")
    srcs = _build.kernel_sources()
    hdrs = sorted((_build.CSRC / "kernels").glob("*.cuh")) + sorted(_build.INCLUDE.glob("*.h"))
    digest = _build._digest(srcs + hdrs, " ".join(_build.NVCC_ARCH + _build.NVCC_FLAGS))
    assert not _build._stale(_build.KERNEL_LIB, digest), \
        "csrc/kernels changed after the last build: run `python -m distributeddataparallel_b200._build`"


def test_async_loss_reader_cpu_values_pass_through_in_order():
    reader = d.AsyncLossReader()
    out = []
    for i in range(5):
        reader.push(torch.tensor(float(i)))
        if reader.pending() > 1:
            out.append(reader.pop())
    out += reader.drain()
    assert out == [0.0, 1.0, 2.0, 3.0, 4.0] and reader.pending() == 0
