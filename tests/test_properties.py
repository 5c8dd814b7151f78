"""Property-based parity checks (hypothesis) of the pieces whose contract is "exactly what torch does":
DistributedSampler index math and the C++ bucket assignment."""
import pytest
import torch
import torch.distributed as dist
from hypothesis import given, settings
from hypothesis import strategies as st
from torch.utils.data import DistributedSampler as Stock

from distributeddataparallel_b200 import DistributedSampler, _C


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@settings(max_examples=150, deadline=None)
@given(n=st.integers(1, 400), ws=st.integers(1, 9), shuffle=st.booleans(), drop_last=st.booleans(),
       seed=st.integers(0, 2 ** 31 - 1), epoch=st.integers(0, 50), data=st.data())
def test_sampler_matches_torch(n, ws, shuffle, drop_last, seed, epoch, data):
    rank = data.draw(st.integers(0, ws - 1))
    a = DistributedSampler(_DS(n), num_replicas=ws, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
    b = Stock(_DS(n), num_replicas=ws, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
    a.set_epoch(epoch)
    b.set_epoch(epoch)
    assert len(a) == len(b) and list(a) == list(b)
    assert a.shard_indices().tolist() == list(b)


_DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.int64]


@settings(max_examples=120, deadline=None)
@given(sizes=st.lists(st.integers(1, 300_000), min_size=1, max_size=40),
       dtype_ids=st.lists(st.integers(0, len(_DTYPES) - 1), min_size=40, max_size=40),
       first_kb=st.integers(1, 2048), cap_kb=st.integers(1, 4096))
def test_bucket_assignment_matches_torch(sizes, dtype_ids, first_kb, cap_kb):
    tensors = [torch.empty(n, dtype=_DTYPES[d]) for n, d in zip(sizes, dtype_ids)]
    limits = [first_kb * 1024, cap_kb * 1024]
    ours, ours_limits = _C.compute_bucket_assignment_by_size(tensors, limits)
    ref, ref_limits = dist._compute_bucket_assignment_by_size(tensors, limits)
    assert [list(b) for b in ours] == [list(b) for b in ref]
    assert list(ours_limits) == list(ref_limits)


@settings(max_examples=60, deadline=None)
@given(sizes=st.lists(st.integers(1, 50_000), min_size=1, max_size=30), buffer_kb=st.integers(1, 512))
def test_coalescing_plan_covers_every_tensor_once_within_the_limit(sizes, buffer_kb):
    tensors = [torch.empty(n, dtype=torch.float32 if i % 3 else torch.int64) for i, n in enumerate(sizes)]
    groups = _C.plan_coalesced_groups(tensors, buffer_kb * 1024)
    flat = [i for g in groups for i in g]
    assert sorted(flat) == list(range(len(tensors)))
    for g in groups:
        assert len({tensors[i].dtype for i in g}) == 1                       # one dtype per message
        nbytes = sum(tensors[i].numel() * tensors[i].element_size() for i in g)
        assert len(g) == 1 or nbytes <= buffer_kb * 1024                      # only a lone oversized tensor may exceed


class _NullComm:
    def launch(self, index, flat, is_last):
        return None


@settings(max_examples=60, deadline=None)
@given(sizes=st.lists(st.integers(1, 60_000), min_size=2, max_size=25),
       thresholds=st.lists(st.integers(1, 400_000), min_size=1, max_size=3),
       first_kb=st.integers(1, 256), cap_kb=st.integers(16, 1024))
def test_tail_split_partitions_the_last_bucket_without_changing_the_reported_layout(sizes, thresholds, first_kb, cap_kb):
    """Reducer.set_tail_splits: launch units are a partition of the parameters in bucket order; every stock bucket but
    the last is one unit; the pieces of the last bucket are contiguous, and every cut sits where the gradients still
    to come first fit under a threshold."""
    params = [torch.nn.Parameter(torch.zeros(n)) for n in sizes]
    alloc = lambda numel, dtype, device: torch.zeros(int(numel), dtype=dtype, device=device)
    red = _C.Reducer(params, [list(range(len(params)))], _NullComm(), alloc, 1, True, False,
                     [f"p{i}" for i in range(len(params))], first_kb * 1024, cap_kb * 1024)
    order = list(reversed(range(len(params))))
    plain = _C.Reducer(params, [list(range(len(params)))], _NullComm(), alloc, 1, True, False,
                       [f"p{i}" for i in range(len(params))], first_kb * 1024, cap_kb * 1024)
    stock = plain.rebuild_buckets(order)
    red.set_tail_splits(thresholds)
    assert [list(b) for b in red.rebuild_buckets(order)] == [list(b) for b in stock]      # reported layout unchanged
    assert list(red.bucket_sizes()) == list(plain.bucket_sizes())
    units = [list(u) for u in red.bucket_params()]
    assert [i for u in units for i in u] == [i for b in stock for i in b]                  # a partition, same order
    head = len(stock) - 1
    assert units[:head] == [list(b) for b in stock[:head]]
    tail_units = units[head:]
    last = list(stock[-1])
    assert [i for u in tail_units for i in u] == last and all(tail_units)
    nbytes = lambda idxs: sum(params[i].numel() * 4 for i in idxs)
    # every piece after the first starts where the remaining bytes first fit under some threshold
    pos = 0
    for u in tail_units[:-1]:
        pos += len(u)
        rest, rest_plus_one = nbytes(last[pos:]), nbytes(last[pos - 1:])
        assert any(rest <= t < rest_plus_one for t in thresholds), (rest, rest_plus_one, thresholds)
    assert sum(red.launch_unit_sizes()) == sum(plain.bucket_sizes())
