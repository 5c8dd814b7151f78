"""Bucket assignment parity with dist._compute_bucket_assignment_by_size."""
import sys

import pytest
import torch
import torch.distributed as dist

import distributeddataparallel_b200 as d
from distributeddataparallel_b200 import _C

MiB = 1 << 20


def _params(model):
    return [p for p in model.parameters() if p.requires_grad]


def test_resnet18_steady_state_buckets():
    ps = _params(d.SimpleCNN())
    assert len(ps) == 62 and sum(p.numel() * 4 for p in ps) == 44_726_568
    rev = list(reversed(ps))
    idx = list(reversed(range(len(ps))))
    buckets, limits = _C.compute_bucket_assignment_by_size(rev, [MiB, 25 * MiB], idx)
    sizes = [sum(ps[i].numel() * 4 for i in b) for b in buckets]
    assert [len(b) for b in buckets] == [5, 15, 42]
    assert sizes == [9_461_800, 26_494_976, 8_769_792]
    ref, ref_limits = dist._compute_bucket_assignment_by_size(rev, [MiB, 25 * MiB], [False] * len(rev), idx)
    assert buckets == ref and limits == ref_limits


def test_first_iteration_single_bucket():
    ps = _params(d.SimpleCNN())
    buckets, _ = _C.compute_bucket_assignment_by_size(ps, [sys.maxsize])
    assert len(buckets) == 1 and len(buckets[0]) == 62


@pytest.mark.parametrize("dtype,expect", [(torch.float32, 5), (torch.bfloat16, 3)])
def test_resnet50_bucket_counts(dtype, expect):
    ps = [p.detach().to(dtype) for p in _params(d.resnet50())]
    rev = list(reversed(ps))
    idx = list(reversed(range(len(ps))))
    buckets, _ = _C.compute_bucket_assignment_by_size(rev, [MiB, 25 * MiB], idx)
    ref, _ = dist._compute_bucket_assignment_by_size(rev, [MiB, 25 * MiB], [False] * len(rev), idx)
    assert buckets == ref and len(buckets) == expect


def test_mixed_dtype_and_sorting_parity():
    g = torch.Generator().manual_seed(0)
    ts = []
    for i in range(40):
        n = int(torch.randint(1, 300_000, (1,), generator=g))
        ts.append(torch.empty(n, dtype=[torch.float32, torch.float16, torch.float64][i % 3]))
    for limits in ([MiB], [100_000, MiB], [1, 2 * MiB, 3 * MiB]):
        mine = _C.compute_bucket_assignment_by_size(ts, limits)
        ref = dist._compute_bucket_assignment_by_size(ts, limits, [False] * len(ts))
        assert mine[0] == ref[0] and mine[1] == ref[1]


def test_coalesced_group_plan():
    ts = [torch.empty(10), torch.empty(5, dtype=torch.int64), torch.empty(10), torch.empty(100)]
    groups = _C.plan_coalesced_groups(ts, 100)
    assert groups == [[0, 2], [1], [3]]


def test_param_metadata_compare():
    a = _C.pack_param_metadata([torch.empty(2, 3), torch.empty(4)])
    b = _C.pack_param_metadata([torch.empty(2, 3), torch.empty(5)])
    assert _C.compare_param_metadata(a, a.clone(), 1) == ""
    assert "size of dim 0" in _C.compare_param_metadata(a, b, 1)


def test_reducer_slots_respect_allocator_storage_offset():
    """Bucket memory may be a view into a bigger arena: slots must land inside that view."""
    import torch.nn as nn

    class Comm:
        launches = 0

        def launch(self, index, flat, is_last):
            class H:
                def wait(self):
                    pass
            return H()

    arena = torch.zeros(10_000)
    cursor = [1000]

    def alloc(numel, dtype, device):
        t = arena[cursor[0]:cursor[0] + numel]
        cursor[0] += numel
        return t

    net = nn.Sequential(nn.Conv2d(3, 4, 3), nn.Linear(4, 2)).to(memory_format=torch.channels_last)
    ps = list(net.parameters())
    red = _C.Reducer(ps, [list(range(len(ps)))], Comm(), alloc, 1)
    lo, hi = arena.data_ptr() + 1000 * 4, arena.data_ptr() + cursor[0] * 4
    for i, p in enumerate(ps):
        slot = red.grad_slot(i)
        assert slot.shape == p.shape and slot.stride() == p.stride()
        assert lo <= slot.data_ptr() and slot.data_ptr() + slot.numel() * 4 <= hi
    red.grad_slot(0).fill_(1.0)
    assert arena[:1000].abs().sum() == 0
    red.detach_hooks()


class _RecordingComm:
    """Minimal comm backend for the reducer: records what is launched."""

    def __init__(self):
        self.calls = []

    def launch(self, index, flat, is_last):
        self.calls.append((index, flat.numel() * flat.element_size(), is_last))

        class _H:
            def wait(self_inner):
                return None

        return _H()


def test_tail_split_streams_the_last_bucket_but_reports_the_stock_layout():
    """set_tail_splits: the last bucket is launched in pieces (cut where the gradients still to come fit under each
    threshold), every piece as soon as its own gradients are ready; bucket_sizes() keeps the stock 5/15/42 layout."""
    model = d.SimpleCNN()
    ps = _params(model)
    comm = _RecordingComm()
    alloc = lambda numel, dtype, device: torch.zeros(int(numel), dtype=dtype, device=device)
    idx, _ = _C.compute_bucket_assignment_by_size(ps, [sys.maxsize])
    red = _C.Reducer(ps, list(reversed(idx)), comm, alloc, 1, True, False, [f"p{i}" for i in range(len(ps))], MiB, 25 * MiB)
    red.set_tail_splits([MiB, 64 << 10])
    order = list(reversed(range(len(ps))))
    red.rebuild_buckets(order)
    assert list(red.bucket_sizes()) == [9_461_800, 26_494_976, 8_769_792]
    units = list(red.launch_unit_sizes())
    # (the exact cut depends on the gradient-ready order; with the real autograd order of this model it is
    # 7 810 048 / 921 600 / 38 144 -- checked on GPUs in test_multigpu.py)
    assert units[:2] == [9_461_800, 26_494_976] and len(units) == 5 and sum(units) == 44_726_568
    assert units[4] <= (64 << 10) < units[3] + units[4] <= MiB < sum(units[2:])
    assert red.num_logical_buckets() == 3 and red.num_buckets() == 5
    # one backward pass: mark gradients ready in order, launches must come in unit order, the last flagged is_last
    red.prepare_for_backward([])
    model(torch.randn(2, 3, 32, 32)).sum().backward()        # hooks fire, the reducer finalises at the end of backward
    assert [c[0] for c in comm.calls] == [0, 1, 2, 3, 4] and [c[2] for c in comm.calls] == [False] * 4 + [True]
    assert red.logging_data()["num_buckets_reduced"] == 3


def test_step_timer_turns_device_stamps_into_logger_figures():
    """C++ timer of the device-timed logger: which reductions belong to the step, busy vs span, exposed communication."""
    T = _C.StepTimer
    marks = [1_000_000, 1_300_000, 1_750_000, 1_800_000]                      # fwd start, bwd start, bwd end, step end (ns)
    stamps = [1_310_000, 1_350_000, 1_500_000, 1_580_000, 1_740_000, 1_770_000, 900_000, 950_000]
    out = T.summarize(marks, stamps, [9_000_000, 26_000_000, 38_144, 1000], 8)
    assert out["forward_us"] == 300.0 and out["backward_us"] == 450.0 and out["step_us"] == 800.0
    assert [u["unit"] for u in out["units"]] == [0, 1, 2]                    # unit 3 carries a stamp older than this backward
    assert out["comm_busy_us"] == 150.0 and out["comm_span_us"] == 460.0
    assert out["exposed_comm_us"] == 20.0 and abs(out["exposed_comm_pct"] - 2.5) < 1e-9
    assert abs(out["units"][1]["bus_gbs"] - 26_000_000 * 1.75 / 80_000) < 1e-6
    # no marks (timing not enabled): per-unit figures only
    bare = T.summarize([0, 0, 0, 0], stamps, [9_000_000, 26_000_000, 38_144, 1000], 2)
    assert "step_us" not in bare and "exposed_comm_us" not in bare and len(bare["units"]) == 4
    t = T(2, 2)                                                               # skip 2, then every 2nd
    kept = [t.record(marks, stamps, [9_000_000, 26_000_000, 38_144, 1000]) for _ in range(7)]
    assert kept == [False, False, True, False, True, False, True] and t.samples() == 3
    avg = t.averages()
    assert avg["avg_exposed_comm_device_time_us"] == 20.0 and abs(avg["avg_exposed_comm_pct"] - 2.5) < 1e-9
