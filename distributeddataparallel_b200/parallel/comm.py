"""Bucket communication backends and the comm-hook surface.

The reducer (csrc/reducer) hands each full gradient bucket to a *comm* object:

    handle = comm.launch(bucket_index, flat_bucket, is_last)      # async
    handle.wait()                                                 # in finalize_backward

Contract: after ``wait()`` the calling stream may read ``flat_bucket`` and it holds the
**average** over ranks (sum * 1/world_size) -- the division the stock reducer applies while
copying gradients in is the backend's job here, so that the P2P kernels can
fuse it into the reduction.

Backends:

* :class:`C10dComm`  -- ``torch.distributed`` collectives (gloo on CPU, NCCL on GPU).  This is
  the *baseline* the reference uses (/root/reference/dpp.py:21,39 -> ProcessGroupNCCL
  all-reduce), kept for CPU plumbing tests, as a cold-path fallback and as the comparison
  arm of the benchmarks.
* :class:`P2PComm` (parallel/p2p.py) -- the product: one-shot / two-shot / NVLS all-reduce
  kernels over NVLink-mapped peer memory, no NCCL call.
* :class:`HookComm` -- adapter for user comm hooks with the ``GradBucket`` API
  (torch ``register_comm_hook``), including built-in ``allreduce_hook``,
  ``fp16_compress_hook`` and ``bf16_compress_hook``.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist


class _Done:
    def wait(self) -> None:
        return None


class _EventHandle:
    """Completion of work issued on another CUDA stream: ``wait()`` orders the current stream after it."""

    def __init__(self, event) -> None:
        self.event = event

    def wait(self) -> None:
        torch.cuda.current_stream().wait_event(self.event)


class _WorkHandle:
    def __init__(self, work, post: Optional[Callable[[], None]] = None) -> None:
        self._work = work
        self._post = post

    def wait(self) -> None:
        if self._work is not None:
            self._work.wait()
        if self._post is not None:
            self._post()


class CommBase:
    name = "base"

    def __init__(self, process_group=None) -> None:
        self.group = process_group
        self.world_size = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self.launches = 0
        # what the summed gradients are divided by: the world size, except inside
        # ``DistributedDataParallel.join(divide_by_initial_world_size=False)`` where it is the number of ranks
        # that still have data this iteration
        self.divisor = self.world_size
        # Optional ``fn(unit_index)`` run right after a launch unit's reduction has been enqueued, ON THE STREAM THAT
        # CARRIES THE REDUCTION: the training step uses it to apply the optimizer update of that unit's parameters
        # while the backward pass is still running (engine.TrainStep, "optimizer in backward").
        self.after_reduce: Optional[Callable[[int], None]] = None
        self._side_stream = None

    # bucket memory -------------------------------------------------------------------
    def allocate(self, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
        return torch.zeros(int(numel), dtype=dtype, device=device)

    # collectives ---------------------------------------------------------------------
    def launch(self, index: int, flat: torch.Tensor, is_last: bool):
        raise NotImplementedError

    def all_reduce_avg_(self, flat: torch.Tensor) -> None:
        self.launch(-1, flat, True).wait()

    def broadcast_(self, flat: torch.Tensor, src: int = 0) -> None:
        if self.world_size > 1:
            dist.broadcast(flat, src=src, group=self.group)

    def barrier(self) -> None:
        if self.world_size > 1:
            dist.barrier(group=self.group)


class C10dComm(CommBase):
    """Baseline data plane: ``dist.all_reduce`` per bucket (what the reference runs)."""

    name = "c10d"

    def launch(self, index: int, flat: torch.Tensor, is_last: bool):
        self.launches += 1
        if flat.is_cuda:
            from ..ops import native

            native.join_side_streams(flat.device)      # weight gradients written on the side streams
        if self.world_size == 1:
            return _Done()
        flat.div_(self.divisor)
        work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return _WorkHandle(work)


class NoComm(CommBase):
    """world_size == 1 without a process group."""

    name = "none"

    def launch(self, index, flat, is_last):
        self.launches += 1
        if not flat.is_cuda:
            return _Done()
        from ..ops import native

        if self.after_reduce is None:
            native.join_side_streams(flat.device)
            return _Done()
        # single process: nothing to reduce, but the unit's optimizer update can already run next to the backward pass
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(device=flat.device)
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(flat.device))
        self._side_stream.wait_event(ready)
        for ev in native.side_stream_events(flat.device):
            self._side_stream.wait_event(ev)
        with torch.cuda.stream(self._side_stream):
            self.after_reduce(index)
        done = torch.cuda.Event()
        done.record(self._side_stream)
        return _EventHandle(done)

    def broadcast_(self, flat, src=0):
        return None

    def barrier(self):
        return None


# ---------------------------------------------------------------------------------------
# comm hooks
# ---------------------------------------------------------------------------------------
class GradBucket:
    """What a comm hook receives: same accessors as ``torch.distributed.GradBucket``
    (T/include/torch/csrc/distributed/c10d/comm.hpp:20-98)."""

    def __init__(self, index: int, buffer: torch.Tensor, gradients: List[torch.Tensor],
                 parameters: List[torch.Tensor], is_last: bool) -> None:
        self._index, self._buffer = index, buffer
        self._gradients, self._parameters, self._is_last = gradients, parameters, is_last

    def index(self) -> int:
        return self._index

    def buffer(self) -> torch.Tensor:
        return self._buffer

    def gradients(self) -> List[torch.Tensor]:
        return self._gradients

    def parameters(self) -> List[torch.Tensor]:
        return self._parameters

    def is_last(self) -> bool:
        return self._is_last

    def set_buffer(self, t: torch.Tensor) -> None:
        self._buffer = t


class HookComm(CommBase):
    """Runs ``hook(state, bucket) -> Future[Tensor]`` for every bucket.  As with stock DDP
    the hook owns the averaging (no pre-division happens before it is called)."""

    name = "hook"

    def __init__(self, process_group, state, hook: Callable, bucket_info: Callable, inner: CommBase) -> None:
        super().__init__(process_group)
        self.state, self.hook, self._bucket_info, self.inner = state, hook, bucket_info, inner

    def allocate(self, numel, dtype, device):
        return self.inner.allocate(numel, dtype, device)

    def broadcast_(self, flat, src=0):
        return self.inner.broadcast_(flat, src)

    def launch(self, index: int, flat: torch.Tensor, is_last: bool):
        self.launches += 1
        # the built-in hooks have fused forms on the P2P data plane: plain average, or fp32 buckets with a 16-bit wire
        # format accumulated in fp32 (one kernel over NVLink instead of cast + NCCL all-reduce + cast back)
        fused = getattr(self.hook, "_b200_fused", None)
        if fused is not None and self.inner.name == "p2p" and self.inner.world_size > 1 and \
                (self.state is None or self.state is self.group or self.state is dist.group.WORLD):
            self.fused_launches = getattr(self, "fused_launches", 0) + 1
            return self.inner.launch(index, flat, is_last, wire=None if fused == "allreduce" else fused)
        if flat.is_cuda:
            from ..ops import native

            native.join_side_streams(flat.device)      # weight gradients written on the side streams
        grads, params = self._bucket_info(index)
        bucket = GradBucket(index, flat, grads, params, is_last)
        fut = self.hook(self.state, bucket)
        if not hasattr(fut, "wait"):
            raise TypeError("comm hook must return a torch.futures.Future")

        def finish() -> None:
            out = fut.wait()
            if isinstance(out, (list, tuple)):
                out = out[0]
            if out.data_ptr() != flat.data_ptr():
                flat.copy_(out.reshape(-1)[: flat.numel()])

        return _WorkHandle(None, finish)


def _group_size(pg) -> int:
    return dist.get_world_size(pg if pg is not None else dist.group.WORLD)


def _allreduce_fut(pg, tensor: torch.Tensor) -> torch.futures.Future:
    pg = pg if pg is not None else dist.group.WORLD
    tensor.div_(_group_size(pg))
    return dist.all_reduce(tensor, group=pg, async_op=True).get_future().then(lambda f: f.value()[0])


def allreduce_hook(process_group, bucket: GradBucket) -> torch.futures.Future:
    """Average the bucket with a plain all-reduce (default_hooks.py:18-50 equivalent)."""
    return _allreduce_fut(process_group, bucket.buffer())


def _compress_hook(dtype: torch.dtype):
    def hook(process_group, bucket: GradBucket) -> torch.futures.Future:
        pg = process_group if process_group is not None else dist.group.WORLD
        buf = bucket.buffer()
        small = buf.to(dtype).div_(_group_size(pg))
        fut = dist.all_reduce(small, group=pg, async_op=True).get_future()

        def back(f):
            buf.copy_(f.value()[0])
            return buf

        return fut.then(back)

    return hook


fp16_compress_hook = _compress_hook(torch.float16)   # default_hooks.py:87-134 equivalent
bf16_compress_hook = _compress_hook(torch.bfloat16)
# markers read by HookComm: on the P2P data plane these hooks are served by fused kernels (csrc/kernels/comm.cu)
allreduce_hook._b200_fused = "allreduce"
fp16_compress_hook._b200_fused = torch.float16
bf16_compress_hook._b200_fused = torch.bfloat16


# ---------------------------------------------------------------------------------------
# coalesced broadcast
# ---------------------------------------------------------------------------------------
def broadcast_coalesced(comm: CommBase, tensors: Sequence[torch.Tensor], buffer_bytes: int, src: int = 0) -> int:
    """Broadcast many tensors from ``src`` in few messages; returns the number of collectives.

    Tensors are grouped per dtype into flat staging buffers of at most ``buffer_bytes``
    (planner in C++: ``_C.plan_coalesced_groups``), each group is one broadcast, then every
    rank scatters the values back.  Used for the constructor-time parameter/buffer sync
    (K2, 250 MiB buffers) and, on the baseline backend, the per-forward BN-buffer sync (K3).
    """
    from .. import _C

    tensors = [t for t in tensors if t.numel() > 0]
    if not tensors or comm.world_size == 1:
        return 0
    groups = _C.plan_coalesced_groups(list(tensors), int(buffer_bytes))
    n = 0
    with torch.no_grad():
        for members in groups:
            group = [tensors[i] for i in members]
            if len(group) == 1 and group[0].is_contiguous():
                comm.broadcast_(group[0], src)
                n += 1
                continue
            flat = torch.cat([t.reshape(-1) for t in group])
            comm.broadcast_(flat, src)
            n += 1
            if comm.rank != src:
                off = 0
                for t in group:
                    t.copy_(flat[off:off + t.numel()].view(t.shape))
                    off += t.numel()
    return n
