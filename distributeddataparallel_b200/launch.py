"""Process launch and process-group bootstrap.

Same names and call shapes as the reference:

* ``setup(rank, world_size)``      -- /root/reference/dpp.py:20-21
* ``cleanup()``                    -- /root/reference/dpp.py:23-24
* ``main()``                       -- /root/reference/dpp.py:60-62 (one process per GPU via
  ``torch.multiprocessing.spawn``)

What is fixed relative to the reference (SURVEY.md §2.1 "quirks"):

* ``setup`` fills in ``MASTER_ADDR`` / ``MASTER_PORT`` when the environment lacks them
  (dpp.py relies on the user having exported them, otherwise env:// rendezvous raises);
* the backend and world size are parameters (dpp.py hard-codes "nccl" and
  ``torch.cuda.device_count()``), so the same code runs on CPU/gloo with world_size 2;
* ``torch.cuda.set_device(rank)`` is called explicitly (dpp.py relies on ``.to(rank)``).

The control plane (rendezvous, TCPStore, process group) stays PyTorch's; only the hot
data plane is replaced by the kernels under ``csrc/kernels``.
"""
from __future__ import annotations

import datetime
import os
import socket
from typing import Callable, Optional, Sequence

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from .config import TrainConfig, config_from_args


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def setup(
    rank: int,
    world_size: int,
    backend: Optional[str] = None,
    master_addr: Optional[str] = None,
    master_port: Optional[int] = None,
    timeout_s: float = 600.0,
) -> None:
    """Join the default process group (dpp.py:20-21), parametrised on backend."""
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    os.environ.setdefault("MASTER_ADDR", master_addr or "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(master_port or 29511))
    if backend == "nccl" and torch.cuda.is_available():
        torch.cuda.set_device(rank % torch.cuda.device_count())
    if dist.is_initialized():
        return
    kwargs = {}
    if backend == "nccl" and torch.cuda.is_available():
        kwargs["device_id"] = torch.device("cuda", rank % torch.cuda.device_count())
    dist.init_process_group(
        backend,
        rank=rank,
        world_size=world_size,
        timeout=datetime.timedelta(seconds=timeout_s),
        **kwargs,
    )


def cleanup() -> None:
    """Leave the process group (dpp.py:23-24)."""
    if dist.is_initialized():
        dist.destroy_process_group()


def spawn(fn: Callable, world_size: int, args: Sequence = (), join: bool = True):
    """``mp.spawn`` with fail-fast semantics (first failing rank tears down the rest and
    its traceback is re-raised in the parent: SURVEY.md §3.5 / §5.3)."""
    return mp.spawn(fn, args=tuple(args), nprocs=world_size, join=join)


def main(argv=None) -> None:
    """CLI entry point: spawn ``train(rank, world_size)`` on every GPU (dpp.py:60-62)."""
    from .train import train  # local import: keep `import package` light

    cfg = config_from_args(argv)
    world_size = cfg.resolved_world_size()
    if "MASTER_PORT" not in os.environ:
        os.environ["MASTER_PORT"] = str(cfg.master_port if cfg.master_port else _free_port())
    os.environ.setdefault("MASTER_ADDR", cfg.master_addr)
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        # launched by torchrun: this process is already one rank
        train(int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), cfg.epochs, cfg)
        return
    if world_size == 1:
        train(0, 1, cfg.epochs, cfg)
        return
    spawn(train, world_size, args=(world_size, cfg.epochs, cfg))


if __name__ == "__main__":
    main()
