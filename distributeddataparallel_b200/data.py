"""Datasets and the device-side batch pipeline.

Reference data path (/root/reference/dpp.py:32-35, :47-48):

    transform = Compose([ToTensor(), Normalize((0.5,), (0.5,))])
    dataset   = CIFAR10(root="data", train=True, download=True, transform=transform)
    loader    = DataLoader(dataset, batch_size=32, sampler=sampler)   # num_workers=0
    ... data.to(rank), target.to(rank)                                # pageable, sync H2D

i.e. per *sample* PIL decode + float conversion on the host, then a synchronous pageable
copy of 393 216 B of fp32 per step.  The B200-first replacement keeps the CIFAR storage
format (uint8 NHWC, 3 072 B / image) all the way to the GPU:

    host:   one index_select of the batch rows into a pinned staging slot   (98 304 B)
    H2D:    cudaMemcpyAsync on a copy stream (4x fewer bytes than the fp32 copy)
    device: one fused kernel: uint8 NHWC -> (x/255 - 0.5)/0.5 fp32 NHWC  (channels_last)

The arithmetic is the same two steps torchvision performs (``ToTensor`` then
``Normalize``), so the values match the reference pipeline bit for bit.

``SyntheticCIFAR10`` has the exact shapes/dtypes of ``torchvision.datasets.CIFAR10``
(``.data`` uint8 [N,32,32,3], ``.targets`` list[int]) because the GPU box has no network
(SURVEY.md §2.1 quirk 3).  ``cifar10()`` returns the real dataset when it is on disk,
downloading on rank 0 only (the reference lets every rank race on the download, quirk 2).
"""
from __future__ import annotations

import os
from typing import Iterator, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist
from torch.utils.data import Dataset

from .sampler import DistributedSampler

CIFAR_MEAN = 0.5   # dpp.py:32 Normalize((0.5,), (0.5,)) broadcast over the 3 channels
CIFAR_STD = 0.5


def normalize_u8(x_u8: torch.Tensor) -> torch.Tensor:
    """ToTensor + Normalize on a uint8 tensor, same op order as torchvision."""
    return (x_u8.to(torch.float32).div(255.0) - CIFAR_MEAN) / CIFAR_STD


class SyntheticCIFAR10(Dataset):
    """Random images with CIFAR-10's storage layout; deterministic in ``seed``."""

    classes = ["airplane", "automobile", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck"]

    def __init__(self, n: int = 50_000, num_classes: int = 10, seed: int = 1234,
                 image_size: int = 32, transform=None) -> None:
        rng = np.random.default_rng(seed)
        self.data = rng.integers(0, 256, size=(n, image_size, image_size, 3), dtype=np.uint8)
        self.targets = rng.integers(0, num_classes, size=(n,), dtype=np.int64).tolist()
        self.transform = transform
        self.num_classes = num_classes

    def __len__(self) -> int:
        return len(self.targets)

    def __getitem__(self, i: int) -> Tuple[torch.Tensor, int]:
        img = torch.from_numpy(self.data[i])                     # HWC uint8
        if self.transform is not None:
            from PIL import Image

            return self.transform(Image.fromarray(self.data[i])), self.targets[i]
        return normalize_u8(img).permute(2, 0, 1).contiguous(), self.targets[i]


def cifar10(root: str = "data", train: bool = True, synthetic: bool = True,
            n: int = 50_000, seed: int = 1234, rank: int = 0) -> Dataset:
    """CIFAR-10 as dpp.py:33 builds it, or the synthetic stand-in."""
    if synthetic:
        return SyntheticCIFAR10(n=n, seed=seed)
    from torchvision import datasets, transforms

    tfm = transforms.Compose([transforms.ToTensor(), transforms.Normalize((CIFAR_MEAN,), (CIFAR_STD,))])
    have = os.path.isdir(os.path.join(root, "cifar-10-batches-py"))
    if dist.is_initialized() and dist.get_world_size() > 1:
        if rank == 0 and not have:
            datasets.CIFAR10(root=root, train=train, download=True)
        dist.barrier()
        return datasets.CIFAR10(root=root, train=train, download=False, transform=tfm)
    return datasets.CIFAR10(root=root, train=train, download=not have, transform=tfm)


class SyntheticTokens(Dataset):
    """Random token streams for the GPT-2 configuration (BASELINE.json config 4)."""

    def __init__(self, n: int = 4096, seq_len: int = 1024, vocab: int = 50257, seed: int = 1234) -> None:
        g = torch.Generator().manual_seed(seed)
        self.tokens = torch.randint(0, vocab, (n, seq_len + 1), generator=g, dtype=torch.int64)

    def __len__(self) -> int:
        return self.tokens.shape[0]

    def __getitem__(self, i: int):
        row = self.tokens[i]
        return row[:-1], row[1:]


class DeviceLoader:
    """Batches straight onto the device from a uint8 image store.

    Equivalent of ``DataLoader(dataset, batch_size=32, sampler=sampler)`` + ``.to(rank)``
    (dpp.py:35,47-48): same batches, same order, ragged last batch kept
    (``drop_last=False``), but staged through pinned memory with the float conversion done
    on the GPU.  Yields ``(data, target)`` with ``data`` fp32 ``[B,3,H,W]`` in channels_last
    memory format (NHWC, what the conv kernels consume) and ``target`` int64 ``[B]``.
    """

    def __init__(self, dataset, batch_size: int, sampler: Optional[DistributedSampler] = None,
                 device: torch.device | str = "cpu", slots: int = 3, drop_last: bool = False) -> None:
        if not hasattr(dataset, "data") or not hasattr(dataset, "targets"):
            raise TypeError("DeviceLoader needs a dataset with uint8 `.data` [N,H,W,C] and `.targets`")
        self.dataset = dataset
        self.batch_size = int(batch_size)
        self.sampler = sampler
        self.device = torch.device(device)
        self.drop_last = drop_last
        data = dataset.data
        self._images = torch.from_numpy(np.ascontiguousarray(data)) if isinstance(data, np.ndarray) else data
        self._labels = torch.as_tensor(dataset.targets, dtype=torch.int64)
        n, h, w, c = self._images.shape
        self._hwc = (h, w, c)
        self._cuda = self.device.type == "cuda"
        self.slots = max(2, slots)
        if self._cuda:
            pin = dict(pin_memory=True)
            self._stage_x = [torch.empty((self.batch_size, h, w, c), dtype=torch.uint8, **pin) for _ in range(self.slots)]
            self._stage_y = [torch.empty((self.batch_size,), dtype=torch.int64, **pin) for _ in range(self.slots)]
            self._dev_x = [torch.empty((self.batch_size, h, w, c), dtype=torch.uint8, device=self.device) for _ in range(self.slots)]
            self._dev_y = [torch.empty((self.batch_size,), dtype=torch.int64, device=self.device) for _ in range(self.slots)]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._ready = [torch.cuda.Event() for _ in range(self.slots)]
            self._consumed = [torch.cuda.Event() for _ in range(self.slots)]
        self.h2d_bytes_per_batch = self.batch_size * (h * w * c + 8)

    def __len__(self) -> int:
        n = len(self.sampler) if self.sampler is not None else len(self._labels)
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def _order(self) -> torch.Tensor:
        if self.sampler is not None:
            return self.sampler.shard_indices()
        return torch.arange(len(self._labels))

    def _stage(self, slot: int, idx: torch.Tensor) -> int:
        b = idx.numel()
        torch.index_select(self._images, 0, idx, out=self._stage_x[slot][:b])
        torch.index_select(self._labels, 0, idx, out=self._stage_y[slot][:b])
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._consumed[slot])
            self._dev_x[slot][:b].copy_(self._stage_x[slot][:b], non_blocking=True)
            self._dev_y[slot][:b].copy_(self._stage_y[slot][:b], non_blocking=True)
            self._ready[slot].record(self._copy_stream)
        return b

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        from .ops import preprocess_u8

        order = self._order()
        nb = len(self)
        bs = self.batch_size
        if not self._cuda:
            for i in range(nb):
                idx = order[i * bs:(i + 1) * bs]
                x = normalize_u8(self._images[idx]).permute(0, 3, 1, 2)   # NHWC storage, NCHW view
                yield x, self._labels[idx]
            return
        # Host-side staging (gather into pinned memory + H2D enqueue on the copy stream) runs in a worker thread
        # that stays up to ``slots - 1`` batches ahead, so the consumer's critical path between two training steps
        # is only: wait-event, the normalise kernel, the label clone.  A slot is re-staged only after the consumer
        # recorded its ``consumed`` event and released the slot.
        import queue
        import threading

        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        free = [threading.Semaphore(1) for _ in range(self.slots)]
        ready_q: "queue.Queue" = queue.Queue()
        stop = threading.Event()

        def worker() -> None:
            try:
                torch.cuda.set_device(dev_index)
                for i in range(nb):
                    slot = i % self.slots
                    free[slot].acquire()
                    if stop.is_set():
                        return
                    if i >= self.slots:
                        self._consumed[slot].synchronize()     # the GPU is done reading this slot's previous batch
                    ready_q.put((i, self._stage(slot, order[i * bs:(i + 1) * bs])))
            except BaseException as exc:                        # surface staging errors in the consumer
                ready_q.put(exc)

        th = threading.Thread(target=worker, name="b200ddp-stage", daemon=True)
        th.start()
        try:
            for i in range(nb):
                item = ready_q.get()
                if isinstance(item, BaseException):
                    raise item
                _, b = item
                slot = i % self.slots
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(self._ready[slot])
                x = preprocess_u8(self._dev_x[slot][:b])          # fused ToTensor+Normalize, NHWC fp32
                y = self._dev_y[slot][:b].clone()
                self._consumed[slot].record(cur)
                free[slot].release()
                yield x, y
        finally:
            stop.set()
            for sem in free:
                sem.release()
