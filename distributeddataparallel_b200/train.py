"""The per-rank training program.

``train(rank, world_size, epochs=5)`` has the reference's name, signature and loop
structure (/root/reference/dpp.py:27-57): setup -> seed 0 -> CIFAR-10 + rank-sharded
sampler + batch 32 -> SimpleCNN -> DDP -> CrossEntropy + SGD(lr=0.01) -> epoch loop with
``sampler.set_epoch`` -> rank-0 print of the local loss every 100 batches (same text) ->
cleanup.  The optional ``cfg`` widens it (backend, synthetic data, checkpointing, graphs).
"""
from __future__ import annotations

import time
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from . import ops
from .checkpoint import load_checkpoint, save_checkpoint
from .config import TrainConfig
from .data import DeviceLoader, SyntheticTokens, cifar10
from .engine import TrainStep
from .launch import cleanup, setup
from .optim import SGD, Adam
from .parallel.ddp import DistributedDataParallel as DDP
from .sampler import DistributedSampler


class CrossEntropyLoss(nn.Module):
    """Mean-reduced softmax cross entropy (dpp.py:40), one fused fwd+bwd kernel on CUDA."""

    def forward(self, logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        return ops.cross_entropy(logits, target)


def build_model(cfg: TrainConfig) -> nn.Module:
    from . import models

    name = cfg.model.lower()
    if name in ("simplecnn", "dpp", "resnet18-cifar"):
        return models.SimpleCNN(cfg.num_classes, pretrained=cfg.pretrained)
    if name == "resnet18":
        return models.resnet18(cfg.num_classes)
    if name == "resnet50":
        return models.resnet50(cfg.num_classes)
    if name.startswith("gpt2"):
        # (not ``models.gpt2(...)``: importing the submodule ``models.gpt2`` rebinds that package attribute)
        from .models.gpt2 import build_gpt2

        return build_gpt2(name, seq_len=cfg.seq_len)
    raise ValueError(f"unknown model {cfg.model!r}")


def build_optimizer(cfg: TrainConfig, params):
    if cfg.optimizer == "adam":
        return Adam(params, lr=cfg.lr, weight_decay=cfg.weight_decay)
    return SGD(params, lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay)   # dpp.py:41


def format_ddp_stats(log: dict, samples: int, seconds: float) -> str:
    """Human-readable summary of ``DistributedDataParallel._get_ddp_logging_data()``: bucket layout,
    collectives issued, host-timed forward / backward / communication shares and -- when the P2P data plane ran with
    ``B200DDP_COMM_TIMING=1`` -- device-timed bucket all-reduce time and bus bandwidth."""
    lines = [f"[ddp] {samples} samples in {seconds:.2f} s = {samples / max(seconds, 1e-9):.0f} samples/s (wall clock)",
             f"[ddp] world {log.get('world_size')} backend {log.get('backend_name')} data plane {log.get('comm_backend')} "
             f"hook {log.get('comm_hook')}",
             f"[ddp] parameters {log.get('num_parameter_tensors')} tensors / {log.get('total_parameter_size_bytes')} B; "
             f"buckets first iteration {log.get('initial_bucket_sizes')} -> rebuilt {log.get('rebuilt_bucket_sizes')}",
             f"[ddp] iterations {log.get('iteration')}, buckets reduced {log.get('num_buckets_reduced')}, "
             f"collectives {log.get('collectives')}"]
    if log.get("avg_backward_compute_time_ms") is not None:
        bwd = float(log.get("avg_backward_compute_time_ms") or 0.0)
        comm = float(log.get("avg_backward_comm_time_ms") or 0.0)
        overlap = float(log.get("avg_backward_compute_comm_overlap_time_ms") or 0.0)
        exposed = max(comm - overlap, 0.0)
        lines.append(f"[ddp] host-timed per iteration: forward {float(log.get('avg_forward_compute_time_ms') or 0):.3f} ms, "
                     f"backward {bwd:.3f} ms, communication {comm:.3f} ms of which {exposed:.3f} ms not overlapped")
    dt = log.get("device_timing") or {}
    if "step_us" in dt:
        lines.append(f"[ddp] device-timed last step: {dt['step_us']:.1f} us = forward {dt.get('forward_us', 0):.1f} + backward "
                     f"{dt.get('backward_us', 0):.1f} + tail; communication busy {dt.get('comm_busy_us', 0):.1f} us over a span of "
                     f"{dt.get('comm_span_us', 0):.1f} us, exposed after backward {dt.get('exposed_comm_us', 0):.1f} us "
                     f"({dt.get('exposed_comm_pct', 0):.1f} % of the step)")
    for u in dt.get("units", []):
        lines.append(f"[ddp]   launch unit {u['unit']}: {u['bytes']} B in {u['us']:.1f} us = {u['bus_gbs']:.0f} GB/s bus")
    cs = log.get("comm_stats") or {}
    if "avg_bucket_allreduce_us" in cs:
        lines.append(f"[ddp] device-timed bucket all-reduce: {cs['avg_bucket_allreduce_us']:.1f} us average over "
                     f"{cs['timed_buckets']} buckets, bus bandwidth {cs['bus_bandwidth_GBs']:.0f} GB/s")
    return "\n".join(lines)


def train(rank: int, world_size: int, epochs: int = 5, cfg: Optional[TrainConfig] = None) -> dict:
    cfg = cfg or TrainConfig()
    device_type = cfg.resolved_device()
    setup(rank, world_size, backend=cfg.resolved_backend(), master_addr=cfg.master_addr,
          master_port=cfg.master_port)                                      # dpp.py:28
    torch.manual_seed(cfg.seed)                                             # dpp.py:29
    device = torch.device("cuda", rank % torch.cuda.device_count()) if device_type == "cuda" else torch.device("cpu")
    ops.set_native(cfg.native_ops)

    is_lm = cfg.model.lower().startswith("gpt2")
    if cfg.dtype == "bf16" and not is_lm:
        raise ValueError("--dtype bf16 applies to the transformer models (gpt2*); the convolutional models run fp32 "
                         "storage with TF32 tensor-core math")
    if is_lm:
        from .models.gpt2 import CONFIGS as _GPT2_CONFIGS

        gcfg = _GPT2_CONFIGS[cfg.model.lower()]
        if cfg.seq_len > gcfg.block_size:
            raise ValueError(f"--seq-len {cfg.seq_len} exceeds the context of {cfg.model} ({gcfg.block_size})")
        dataset = SyntheticTokens(n=cfg.dataset_size, seq_len=cfg.seq_len, vocab=gcfg.vocab_size, seed=1234)
    elif cfg.image_size != 32:
        from .data import SyntheticCIFAR10

        if not cfg.synthetic:
            raise ValueError("--image-size other than 32 needs --synthetic (CIFAR-10 images are 32x32)")
        dataset = SyntheticCIFAR10(n=cfg.dataset_size, seed=1234, image_size=cfg.image_size)
    else:
        dataset = cifar10(cfg.data_root, train=True, synthetic=cfg.synthetic, n=cfg.dataset_size, rank=rank)   # :33
    sampler = DistributedSampler(dataset, num_replicas=world_size, rank=rank, shuffle=True)               # :34
    if hasattr(dataset, "data"):
        dataloader = DeviceLoader(dataset, cfg.batch_size, sampler, device)                               # :35
    else:
        from torch.utils.data import DataLoader

        dataloader = DataLoader(dataset, batch_size=cfg.batch_size, sampler=sampler)

    model = build_model(cfg).to(device)                                     # dpp.py:38
    if cfg.dtype == "bf16":
        model = model.to(torch.bfloat16)
    model = DDP(model, device_ids=[device.index] if device.type == "cuda" else None,
                broadcast_buffers=cfg.broadcast_buffers, bucket_cap_mb=cfg.bucket_cap_mb,
                first_bucket_mb=cfg.first_bucket_mb, comm=cfg.comm, allreduce_algo=cfg.allreduce_algo)   # :39
    if cfg.print_ddp_stats:
        model.enable_device_timing()          # before the step is captured: the marks become graph nodes
    criterion = CrossEntropyLoss()                                          # dpp.py:40
    optimizer = build_optimizer(cfg, model.parameters())                    # dpp.py:41

    start_epoch, global_step = 0, 0
    if cfg.resume:
        info = load_checkpoint(cfg.resume, model, optimizer, sampler, map_location=device)
        start_epoch, global_step = info["epoch"], info["step"]

    step_fn = TrainStep(model, criterion, optimizer, cuda_graph=cfg.cuda_graph)
    history = []
    t0 = time.time()
    seen = 0
    done = False
    for epoch in range(start_epoch, epochs):                                # dpp.py:44
        model.train()                                                       # dpp.py:45
        sampler.set_epoch(epoch)                                            # dpp.py:46
        for batch_idx, (data, target) in enumerate(dataloader):             # dpp.py:47
            if data.device != device:
                data, target = data.to(device), target.to(device)           # dpp.py:48
            loss = step_fn(data, target)                                    # dpp.py:49-53
            global_step += 1
            seen += data.shape[0]
            if batch_idx % cfg.log_interval == 0:
                model.check_health(sync=False)      # a collective that timed out on this rank raises here
            if batch_idx % cfg.log_interval == 0 and rank == 0:             # dpp.py:54-55
                val = loss.item()
                history.append(val)
                line = f"Epoch {epoch}, Batch {batch_idx}, Loss: {val}"
                if cfg.print_throughput:
                    line += f"  [{seen * world_size / max(time.time() - t0, 1e-9):.0f} samples/s]"
                print(line, flush=True)
            if cfg.max_steps is not None and global_step >= cfg.max_steps:
                done = True
                break
        if cfg.checkpoint_path and (done or epoch == epochs - 1 or
                                    (cfg.checkpoint_every and (epoch + 1) % cfg.checkpoint_every == 0)):
            model.check_health(sync=True)           # never persist parameters updated from a failed reduction
            save_checkpoint(cfg.checkpoint_path, model, optimizer, epoch=epoch + 1, step=global_step, sampler=sampler)
        if done:
            break
    stats = {"steps": global_step, "loss_history": history, "graph_replays": step_fn.replays,
             "ddp": model._get_ddp_logging_data()}
    if cfg.print_ddp_stats and device.type == "cuda":
        stats["ddp"]["device_timing"] = model.device_timing()
    if cfg.print_ddp_stats and rank == 0:
        print(format_ddp_stats(stats["ddp"], seen * world_size, time.time() - t0), flush=True)
    del step_fn
    model.check_health(sync=True)
    cleanup()                                                               # dpp.py:57
    return stats
