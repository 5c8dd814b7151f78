"""Rank-sharded epoch sampler.

Behavioural contract (index-for-index) is the sampler the reference builds at
/root/reference/dpp.py:34 and re-seeds at dpp.py:46:

    DistributedSampler(dataset, num_replicas=world_size, rank=rank, shuffle=True)
    sampler.set_epoch(epoch)

i.e. a seeded global permutation (seed + epoch), wrapped-around padding so that every
rank gets the same number of samples, and a *strided* (not blocked) split
``perm[rank::world]``.  SURVEY.md §3.4 has the derivation; tests/test_sampler.py checks
parity against ``torch.utils.data.DistributedSampler`` over a parameter grid.

Differences by design: the shard is materialised as one int64 tensor
(:meth:`shard_indices`) so the device loader can gather a whole epoch's batches with
tensor indexing instead of iterating Python ints.
"""
from __future__ import annotations

from typing import Iterator, Optional, Sized

import torch
import torch.distributed as dist
from torch.utils.data import Sampler


class DistributedSampler(Sampler[int]):
    def __init__(
        self,
        dataset: Sized,
        num_replicas: Optional[int] = None,
        rank: Optional[int] = None,
        shuffle: bool = True,
        seed: int = 0,
        drop_last: bool = False,
    ) -> None:
        if num_replicas is None:
            num_replicas = dist.get_world_size() if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_initialized() else 0
        if not 0 <= rank < num_replicas:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = int(num_replicas)
        self.rank = int(rank)
        self.shuffle = bool(shuffle)
        self.seed = int(seed)
        self.drop_last = bool(drop_last)
        self.epoch = 0
        n = len(dataset)
        if self.drop_last and n % self.num_replicas:
            # largest per-rank count such that per_rank * world <= n
            self.num_samples = -(-(n - self.num_replicas) // self.num_replicas)
        else:
            self.num_samples = -(-n // self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    # -- index math -----------------------------------------------------------------
    def global_order(self) -> torch.Tensor:
        """The padded / truncated global ordering shared by every rank this epoch."""
        n = len(self.dataset)
        if self.shuffle:
            gen = torch.Generator()
            gen.manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=gen)
        else:
            order = torch.arange(n)
        if self.drop_last:
            return order[: self.total_size]
        short = self.total_size - n
        if short > 0:
            reps = -(-self.total_size // n)
            order = order.repeat(reps)[: self.total_size] if n else order
        return order

    def shard_indices(self) -> torch.Tensor:
        """This rank's indices for the current epoch as an int64 tensor."""
        shard = self.global_order()[self.rank :: self.num_replicas]
        assert shard.numel() == self.num_samples
        return shard

    def __iter__(self) -> Iterator[int]:
        return iter(self.shard_indices().tolist())

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    # -- resume support (sampler resume == set_epoch) --------------------
    def state_dict(self) -> dict:
        return {"epoch": self.epoch, "seed": self.seed}

    def load_state_dict(self, state: dict) -> None:
        self.epoch = int(state.get("epoch", 0))
        self.seed = int(state.get("seed", self.seed))
