"""Configuration: the reference's literals as defaults, everything overridable.

The reference (/root/reference/dpp.py) has no flag system at all; its
constants are literals: backend "nccl" (dpp.py:21), epochs=5 (:27), batch_size=32 (:35),
lr=0.01 (:41), log interval 100 (:54), data root "data" (:33), 10 classes (:15), seed 0
(:29).  ``TrainConfig()`` reproduces exactly those; CLI flags / env vars widen them
(backend, world size, synthetic data, dtype, bucket cap, all-reduce algorithm, graphs).
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass, field
from typing import Optional


@dataclass
class TrainConfig:
    # --- reference literals -------------------------------------------------------
    backend: str = "nccl"              # dpp.py:21
    epochs: int = 5                    # dpp.py:27
    batch_size: int = 32               # dpp.py:35 (per rank)
    lr: float = 0.01                   # dpp.py:41
    momentum: float = 0.0
    weight_decay: float = 0.0
    log_interval: int = 100            # dpp.py:54
    data_root: str = "data"            # dpp.py:33
    num_classes: int = 10              # dpp.py:15
    seed: int = 0                      # dpp.py:29
    model: str = "simplecnn"           # dpp.py:11-18  (resnet18 | resnet50 | gpt2-medium | ...)
    # --- additions ----------------------------------------------------------------
    world_size: Optional[int] = None   # default: all visible GPUs (dpp.py:61) or 1 on CPU
    device: str = "auto"               # auto | cuda | cpu
    synthetic: bool = True             # the GPU box has no network: synthetic CIFAR-shaped data
    dataset_size: int = 50_000         # CIFAR-10 train split size
    pretrained: bool = False           # dpp.py:14 asks for ImageNet weights; needs network
    dtype: str = "fp32"                # fp32 (tf32 tensor-core math) | bf16
    optimizer: str = "sgd"             # sgd | adam
    bucket_cap_mb: float = 25.0        # stock DDP default
    first_bucket_mb: float = 1.0
    allreduce_algo: str = "auto"       # auto | oneshot | twoshot | nvls | push  (P2P data plane; NCCL is --comm c10d)
    comm: str = "auto"                 # auto | p2p (own NVLink kernels) | c10d (NCCL/gloo baseline)
    cuda_graph: bool = True            # capture the whole step once shapes/buckets are final
    native_ops: bool = True            # route model math through the sm_100a kernels
    broadcast_buffers: bool = True
    max_steps: Optional[int] = None    # stop after this many optimizer steps (tests / bench)
    checkpoint_path: Optional[str] = None
    checkpoint_every: int = 0          # epochs; 0 = only at the end when a path is given
    resume: Optional[str] = None
    master_addr: str = "127.0.0.1"
    master_port: int = 29511
    seq_len: int = 1024                # GPT-2 only
    image_size: int = 32               # 32 for CIFAR shaped, 224 for ImageNet shaped
    print_throughput: bool = False
    print_ddp_stats: bool = False      # rank 0 prints the reducer / data-plane summary when training ends
    extra: dict = field(default_factory=dict)

    def resolved_device(self) -> str:
        import torch

        if self.device == "auto":
            return "cuda" if torch.cuda.is_available() else "cpu"
        return self.device

    def resolved_backend(self) -> str:
        dev = self.resolved_device()
        if dev == "cpu" and self.backend == "nccl":
            return "gloo"
        return self.backend

    def resolved_world_size(self) -> int:
        import torch

        if self.world_size:
            return int(self.world_size)
        if self.resolved_device() == "cuda":
            return max(1, torch.cuda.device_count())   # dpp.py:61
        return 1

    def replace(self, **kw) -> "TrainConfig":
        return dataclasses.replace(self, **kw)


def _add_bool(p: argparse.ArgumentParser, name: str, default: bool, help: str) -> None:
    dest = name.replace("-", "_")
    p.add_argument(f"--{name}", dest=dest, action="store_true", default=default, help=help)
    p.add_argument(f"--no-{name}", dest=dest, action="store_false")


def build_parser() -> argparse.ArgumentParser:
    d = TrainConfig()
    p = argparse.ArgumentParser(
        prog="distributeddataparallel_b200",
        description="B200-native data-parallel trainer (reference: dpp.py)",
    )
    p.add_argument("--backend", default=d.backend)
    p.add_argument("--world-size", type=int, default=None)
    p.add_argument("--device", default=d.device, choices=["auto", "cuda", "cpu"])
    p.add_argument("--epochs", type=int, default=d.epochs)
    p.add_argument("--batch-size", type=int, default=d.batch_size)
    p.add_argument("--lr", type=float, default=d.lr)
    p.add_argument("--momentum", type=float, default=d.momentum)
    p.add_argument("--weight-decay", type=float, default=d.weight_decay)
    p.add_argument("--optimizer", default=d.optimizer, choices=["sgd", "adam"])
    p.add_argument("--log-interval", type=int, default=d.log_interval)
    p.add_argument("--data-root", default=d.data_root)
    p.add_argument("--seed", type=int, default=d.seed)
    p.add_argument("--model", default=d.model)
    p.add_argument("--dataset-size", type=int, default=d.dataset_size)
    p.add_argument("--dtype", default=d.dtype, choices=["fp32", "bf16"])
    p.add_argument("--bucket-cap-mb", type=float, default=d.bucket_cap_mb)
    p.add_argument("--allreduce-algo", default=d.allreduce_algo,
                   choices=["auto", "oneshot", "twoshot", "nvls", "push"])
    p.add_argument("--comm", default=d.comm, choices=["auto", "p2p", "c10d"])
    p.add_argument("--max-steps", type=int, default=None)
    p.add_argument("--checkpoint-path", default=None)
    p.add_argument("--checkpoint-every", type=int, default=0)
    p.add_argument("--resume", default=None)
    p.add_argument("--master-addr", default=os.environ.get("MASTER_ADDR", d.master_addr))
    p.add_argument("--master-port", type=int, default=int(os.environ.get("MASTER_PORT", d.master_port)))
    p.add_argument("--seq-len", type=int, default=d.seq_len)
    p.add_argument("--image-size", type=int, default=d.image_size)
    _add_bool(p, "synthetic", d.synthetic, "synthetic CIFAR-shaped data (no network needed)")
    _add_bool(p, "pretrained", d.pretrained, "load ImageNet weights as dpp.py:14 does (needs network)")
    _add_bool(p, "cuda-graph", d.cuda_graph, "capture the whole training step in a CUDA graph")
    _add_bool(p, "native-ops", d.native_ops, "run the model on the sm_100a kernels")
    _add_bool(p, "broadcast-buffers", d.broadcast_buffers, "re-broadcast BN buffers from rank 0 each forward")
    _add_bool(p, "print-throughput", d.print_throughput, "print samples/s next to the loss line")
    _add_bool(p, "print-ddp-stats", d.print_ddp_stats,
              "rank 0 prints bucket sizes, collective counts and (with B200DDP_COMM_TIMING=1) per-bucket "
              "all-reduce time / bus GB/s when training ends")
    return p


def config_from_args(argv=None) -> TrainConfig:
    ns = build_parser().parse_args(argv)
    fields = {f.name for f in dataclasses.fields(TrainConfig)}
    return TrainConfig(**{k: v for k, v in vars(ns).items() if k in fields})
