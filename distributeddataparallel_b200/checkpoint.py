"""Rank-0 checkpoint save / load.

The reference has no checkpointing at all; the contract implemented
here is the one its objects imply: the model entry is exactly
``DDP(SimpleCNN()).state_dict()`` -- 122 keys, every key ``module.model.<resnet18 key>``,
102 fp32 + 20 int64 tensors -- and the optimizer entry is ``optim.SGD.state_dict()``.  Files
written here load into a stock ``torch.nn.parallel.DistributedDataParallel(SimpleCNN())``
and vice versa (tests/test_checkpoint.py).

Layout of the file (``torch.save``): ``{"model", "optimizer", "epoch", "step", "sampler",
"meta"}``.  Tensors are saved in standard contiguous layout regardless of the
channels_last storage used at run time.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist

FORMAT_VERSION = 1


def _rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def model_state(model: torch.nn.Module) -> Dict[str, torch.Tensor]:
    sd = model.state_dict()
    out = type(sd)()
    for k, v in sd.items():
        out[k] = v.detach().to("cpu").contiguous() if isinstance(v, torch.Tensor) else v
    if hasattr(sd, "_metadata"):
        out._metadata = sd._metadata
    return out


def save_checkpoint(path: str, model: torch.nn.Module, optimizer: Optional[torch.optim.Optimizer] = None,
                    epoch: int = 0, step: int = 0, sampler=None, extra: Optional[dict] = None,
                    barrier: bool = True) -> None:
    """Rank 0 writes atomically (tmp + rename); the other ranks only join the barrier."""
    if _rank() == 0:
        payload: Dict[str, Any] = {
            "model": model_state(model),
            "optimizer": optimizer.state_dict() if optimizer is not None else None,
            "epoch": int(epoch),
            "step": int(step),
            "sampler": sampler.state_dict() if hasattr(sampler, "state_dict") else None,
            "meta": {"format": FORMAT_VERSION, "world_size": dist.get_world_size() if dist.is_initialized() else 1,
                     **(extra or {})},
        }
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = f"{path}.tmp.{os.getpid()}"
        torch.save(payload, tmp)
        os.replace(tmp, path)
    if barrier and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def load_checkpoint(path: str, model: torch.nn.Module, optimizer: Optional[torch.optim.Optimizer] = None,
                    sampler=None, map_location="cpu", strict: bool = True, rank0_only_read: bool = False) -> dict:
    """Restore model/optimizer/sampler.  With ``rank0_only_read`` only rank 0 touches the file
    and the values reach the other ranks through the wrapper's broadcast path."""
    ws = dist.get_world_size() if dist.is_initialized() else 1
    payload = None
    if not rank0_only_read or _rank() == 0:
        payload = torch.load(path, map_location=map_location, weights_only=False)
    if rank0_only_read and ws > 1:
        box = [payload if _rank() == 0 else None]
        dist.broadcast_object_list(box, src=0)
        payload = box[0]
    sd = payload["model"] if "model" in payload else payload
    want_prefix = any(k.startswith("module.") for k in model.state_dict().keys())
    have_prefix = any(k.startswith("module.") for k in sd.keys())
    if want_prefix and not have_prefix:
        sd = {f"module.{k}": v for k, v in sd.items()}
    elif have_prefix and not want_prefix:
        sd = {k[len("module."):]: v for k, v in sd.items()}
    model.load_state_dict(sd, strict=strict)
    if optimizer is not None and payload.get("optimizer") is not None:
        optimizer.load_state_dict(payload["optimizer"])
    if sampler is not None and payload.get("sampler") is not None and hasattr(sampler, "load_state_dict"):
        sampler.load_state_dict(payload["sampler"])
    return {"epoch": payload.get("epoch", 0), "step": payload.get("step", 0), "meta": payload.get("meta", {})}
