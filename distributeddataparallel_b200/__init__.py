"""distributeddataparallel_b200 -- a Blackwell (B200, sm_100a) native data-parallel trainer.

Public surface mirrors the reference script /root/reference/dpp.py:

    SimpleCNN, setup(rank, world_size), cleanup(), train(rank, world_size, epochs=5), main()

plus the pieces that script imports from torch / torchvision, re-implemented here:
``DistributedDataParallel`` (alias ``DDP``), ``DistributedSampler``, ``SGD`` / ``Adam``,
``CrossEntropyLoss``, the ResNet family, and checkpoint helpers.
"""
from __future__ import annotations

__version__ = "0.1.0"


def _load_host_extension():
    import importlib

    try:
        return importlib.import_module("distributeddataparallel_b200._C")
    except ImportError:
        from . import _build

        _build.build_host_extension()
        return importlib.import_module("distributeddataparallel_b200._C")


import torch as _torch  # noqa: E402  (must be loaded before the extension for symbol resolution)

_C = _load_host_extension()

from .config import TrainConfig, config_from_args  # noqa: E402
from .sampler import DistributedSampler  # noqa: E402
from .launch import setup, cleanup, main, spawn  # noqa: E402
from .parallel.ddp import DistributedDataParallel  # noqa: E402
from .parallel import comm as comm_hooks  # noqa: E402
from .optim import SGD, Adam  # noqa: E402
from .models import SimpleCNN, resnet18, resnet34, resnet50  # noqa: E402
from .train import train, CrossEntropyLoss, build_model  # noqa: E402
from .engine import TrainStep  # noqa: E402
from .checkpoint import save_checkpoint, load_checkpoint  # noqa: E402
from .data import SyntheticCIFAR10, DeviceLoader, cifar10  # noqa: E402
from .utils.loss_reader import AsyncLossReader  # noqa: E402

DDP = DistributedDataParallel

__all__ = [
    "TrainConfig", "config_from_args", "DistributedSampler", "setup", "cleanup", "main", "spawn",
    "DistributedDataParallel", "DDP", "SGD", "Adam", "SimpleCNN", "resnet18", "resnet34", "resnet50",
    "train", "CrossEntropyLoss", "build_model", "TrainStep", "save_checkpoint", "load_checkpoint",
    "SyntheticCIFAR10", "DeviceLoader", "cifar10", "comm_hooks", "AsyncLossReader",
]
