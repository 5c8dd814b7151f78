"""Index-for-index parity with torch.utils.data.DistributedSampler."""
import itertools

import pytest
import torch
from torch.utils.data import DistributedSampler as Stock

from distributeddataparallel_b200 import DistributedSampler


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("n,ws", [(50, 1), (50, 2), (51, 2), (10, 4), (3, 8), (1, 3), (1000, 8), (17, 5)])
@pytest.mark.parametrize("shuffle", [True, False])
@pytest.mark.parametrize("drop_last", [True, False])
def test_parity_grid(n, ws, shuffle, drop_last):
    for rank, epoch, seed in itertools.product(range(ws), (0, 3), (0, 7)):
        a = DistributedSampler(_DS(n), num_replicas=ws, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        b = Stock(_DS(n), num_replicas=ws, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        a.set_epoch(epoch)
        b.set_epoch(epoch)
        assert len(a) == len(b)
        assert list(a) == list(b)


def test_cifar_sized_shards_cover_everything():
    n = 50_000
    for ws in (1, 2, 4, 8):
        seen = torch.cat([DistributedSampler(_DS(n), ws, r).shard_indices() for r in range(ws)])
        assert seen.numel() == n and seen.unique().numel() == n
    # batches per rank per epoch and ragged tail from SURVEY §3.4
    for ws, (nb, tail) in {1: (1563, 16), 2: (782, 8), 4: (391, 20), 8: (196, 10)}.items():
        per = len(DistributedSampler(_DS(n), ws, 0))
        assert -(-per // 32) == nb and per - (nb - 1) * 32 == tail


def test_bad_rank():
    with pytest.raises(ValueError):
        DistributedSampler(_DS(4), num_replicas=2, rank=2)


def test_state_roundtrip():
    s = DistributedSampler(_DS(10), 2, 0)
    s.set_epoch(4)
    t = DistributedSampler(_DS(10), 2, 0)
    t.load_state_dict(s.state_dict())
    assert list(s) == list(t)
