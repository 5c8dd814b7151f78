"""``DistributedDataParallel``: the data-parallel module wrapper.

Drop-in for the object the reference creates at /root/reference/dpp.py:39

    model = DDP(model, device_ids=[rank])

re-implemented for the B200 data plane (this is *not* a wrapper around
``torch.nn.parallel.DistributedDataParallel``).  Observable behaviour kept from the stock
wrapper (citing T/nn/parallel/distributed.py):

* constructor: parameter shape verification across ranks (K1), then rank 0's parameters
  and buffers are broadcast to everyone (K2);
* ``self.module`` holds the user model, so ``state_dict()`` keys carry the ``module.``
  prefix (the checkpoint contract);
* iteration 0 reduces all gradients as one bucket; before iteration 1 the buckets are
  rebuilt once in gradient-ready order with [1 MiB, 25 MiB] limits, using rank 0's order
  (K5) -- for the dpp.py model that yields 5/15/42 parameters, 9 461 800 / 26 494 976 /
  8 769 792 bytes;
* every forward re-broadcasts module buffers (BN running stats) from rank 0 (K3);
* gradients are averaged over ranks; ``no_sync()``, ``register_comm_hook()``,
  ``find_unused_parameters`` and ``_get_ddp_logging_data()`` exist with the same meaning.

What is different by design:

* gradients live inside the flat buckets (``gradient_as_bucket_view=True`` is the default);
  our weight-gradient kernels write straight into the bucket slot a parameter exposes as
  ``param._grad_slot``, removing the 124 copy kernels per step of the stock reducer;
* the bucket reduction is a fused (sum, x 1/world, cast) P2P/NVLS kernel over NVLink-mapped
  memory (``comm="p2p"``); ``comm="c10d"`` keeps NCCL/gloo as the baseline;
* the reducer itself is native C++ (csrc/reducer), driven by grad-accumulator post hooks.
"""
from __future__ import annotations

import contextlib
import os
import sys
import weakref
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn

from .. import _C
from .comm import C10dComm, CommBase, HookComm, NoComm, broadcast_coalesced

MiB = 1 << 20


class DistributedDataParallel(nn.Module):
    def __init__(
        self,
        module: nn.Module,
        device_ids: Optional[Sequence] = None,
        output_device=None,
        dim: int = 0,
        broadcast_buffers: bool = True,
        init_sync: bool = True,
        process_group=None,
        bucket_cap_mb: Optional[float] = None,
        find_unused_parameters: bool = False,
        gradient_as_bucket_view: bool = True,
        static_graph: bool = False,
        comm: str | CommBase = "auto",
        first_bucket_mb: float = 1.0,
        allreduce_algo: str = "auto",
        tail_splits: Optional[Sequence[int]] = None,
    ) -> None:
        super().__init__()
        named = [(n, p) for n, p in self._named_params(module)]
        if not named:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any "
                               "parameter that requires a gradient.")
        self._param_names = [n for n, _ in named]
        self._params = [p for _, p in named]
        devices = {p.device for p in self._params}
        if len(devices) != 1:
            raise ValueError(f"all parameters must be on one device, found {sorted(map(str, devices))}")
        self.device = next(iter(devices))
        if device_ids is not None and len(device_ids) > 1:
            raise ValueError("device_ids can only be None or contain a single element")
        if device_ids is not None and self.device.type == "cuda":
            want = torch.device("cuda", _device_index(device_ids[0]))
            if want != self.device:
                raise ValueError(f"device_ids={list(device_ids)} but the module is on {self.device}")
        self.device_ids = [self.device.index] if self.device.type == "cuda" else None
        self.output_device = self.device_ids[0] if self.device_ids else None
        self.dim = dim
        self.module = module                      # origin of the "module." key prefix
        self.process_group = process_group
        self.broadcast_buffers = broadcast_buffers
        self.find_unused_parameters = find_unused_parameters
        self.gradient_as_bucket_view = gradient_as_bucket_view
        self.static_graph = static_graph
        self.require_backward_grad_sync = True
        self.bucket_bytes_cap = int((25.0 if bucket_cap_mb is None else bucket_cap_mb) * MiB)
        self.first_bucket_bytes = int(first_bucket_mb * MiB)
        self.broadcast_bucket_size = 250 * MiB
        self.allreduce_algo = allreduce_algo
        self._comm_hook_name = "builtin"
        self._collectives = {"broadcast": 0, "allgather": 0, "allreduce_control": 0}
        self._join_enabled = False
        self._join_iterations = 0
        self._buffer_src = 0               # rank whose buffers are broadcast before each forward
        self._join_divide_by_world = True
        self._join_throw = False
        self._rebuilt_bucket_sizes: List[int] = []
        self._graph_mode = False
        self._buffers_ahead = False        # rank 0's buffers for the next forward were already broadcast
        self._timing = False               # device-timed logger marks (enable_device_timing)
        self._marks = None
        self._step_timer = None            # C++ running averages over the steps device_timing() was called for

        self.world_size = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self._comm = self._make_comm(comm)

        if init_sync and self.world_size > 1:
            self._verify_params_across_processes()
            self._sync_module_states()
        ignored = set(getattr(module, "_ddp_params_and_buffers_to_ignore", ()) or ())
        self._buffers_for_sync = [b for n, b in module.named_buffers() if n not in ignored]

        # Iteration 0: every gradient in ONE bucket (size limit "infinite") unless unused
        # parameters must be searched, exactly like the stock wrapper (distributed.py:1224).
        if find_unused_parameters and not static_graph:
            limits = [self.first_bucket_bytes, self.bucket_bytes_cap]
        else:
            limits = [sys.maxsize]
        indices, _ = _C.compute_bucket_assignment_by_size(self._params, limits)
        self.reducer = _C.Reducer(
            self._params, list(reversed(indices)), self._comm, self._comm.allocate, self.world_size,
            gradient_as_bucket_view, find_unused_parameters, self._param_names,
            self.first_bucket_bytes, self.bucket_bytes_cap)
        self._initial_bucket_sizes = list(self.reducer.bucket_sizes())
        self._publish_grad_slots()
        me = weakref.ref(self)
        self.reducer.set_backward_end_callback(lambda: me() is not None and me()._mark(2))
        if os.environ.get("B200DDP_DEVICE_TIMING", "0") != "0":
            self.enable_device_timing()
        if tail_splits is None and self._comm.name == "p2p" and self.world_size > 1:
            # stream the last bucket (the one whose final gradient is the stem's weight gradient): pieces are
            # reduced as they become ready, the last few KB go through a small-message kernel.  On by default on
            # the P2P data plane (each extra collective costs ~10 us there; over NCCL / gloo it would not pay);
            # ``tail_splits=[...]`` byte thresholds from the end of the bucket force it on any backend, ``[]`` disables.
            spec = os.environ.get("B200DDP_TAIL_SPLITS", f"{1 << 20},{64 << 10}")
            tail_splits = [int(v) for v in spec.split(",") if v.strip()]
        if tail_splits:
            self.reducer.set_tail_splits([int(v) for v in tail_splits])

    # -- construction helpers --------------------------------------------------------------
    @staticmethod
    def _named_params(module: nn.Module):
        # stock contract (distributed.py:732-736, 2331): names listed in ``module._ddp_params_and_buffers_to_ignore``
        # are neither reduced nor broadcast
        ignored = set(getattr(module, "_ddp_params_and_buffers_to_ignore", ()) or ())
        seen = set()
        for mod_name, mod in module.named_modules():
            for pname, p in mod.named_parameters(recurse=False):
                fqn = f"{mod_name}.{pname}" if mod_name else pname
                if p.requires_grad and id(p) not in seen and fqn not in ignored:
                    seen.add(id(p))
                    yield fqn, p

    @staticmethod
    def _set_params_and_buffers_to_ignore_for_model(module: nn.Module, params_and_buffers_to_ignore) -> None:
        """Call before wrapping: fully-qualified parameter / buffer names DDP must leave alone."""
        module._ddp_params_and_buffers_to_ignore = list(params_and_buffers_to_ignore)

    def _set_static_graph(self) -> None:
        """Accepted for API compatibility: the whole-step CUDA graph (engine.TrainStep) is this framework's
        static-graph mode; the eager reducer does not need the hint."""
        self.static_graph = True

    def will_sync_module_buffers(self) -> bool:
        return (self.require_backward_grad_sync and self.broadcast_buffers and self.world_size > 1
                and len(self._buffers_for_sync) > 0)

    def _make_comm(self, comm) -> CommBase:
        if isinstance(comm, CommBase):
            return comm
        if self.world_size == 1 and not dist.is_initialized():
            return NoComm()
        auto = comm == "auto"
        if auto:
            comm = "p2p" if self.device.type == "cuda" and self.world_size > 1 else "c10d"
        if comm == "p2p":
            from .p2p import P2PComm

            grad_bytes = sum(p.numel() * p.element_size() for p in self._params)
            buf_bytes = sum(b.numel() * b.element_size() + 16 for b in self.module.buffers())
            from .p2p import ArenaUnavailable

            p2p, ok = None, 1
            try:
                p2p = P2PComm(self.process_group, self.device, algo=self.allreduce_algo)
                p2p.reserve(grad_bytes, buf_bytes)
            except ArenaUnavailable as e:
                # raised by EVERY rank at the same point of the staged arena set-up (parallel/p2p.py), so all
                # ranks fall back (or fail) together and no peer is left inside a collective
                if p2p is not None:
                    p2p.close()
                if not auto:
                    raise
                ok, self._p2p_error = 0, str(e)
            if ok:
                return p2p
            import warnings

            warnings.warn("P2P arena unavailable (%s); falling back to the c10d backend"
                          % getattr(self, "_p2p_error", "a peer failed"))
            return C10dComm(self.process_group)
        if comm in ("c10d", "nccl", "gloo"):
            return C10dComm(self.process_group)
        raise ValueError(f"unknown comm backend {comm!r}")

    def _verify_params_across_processes(self) -> None:
        mine = _C.pack_param_metadata(self._params)
        dev = self.device if dist.get_backend(self.process_group) == "nccl" else torch.device("cpu")
        n = torch.tensor([mine.numel()], dtype=torch.int64, device=dev)
        sizes = [torch.zeros_like(n) for _ in range(self.world_size)]
        dist.all_gather(sizes, n, group=self.process_group)
        self._collectives["allgather"] += 1
        width = max(int(s.item()) for s in sizes)
        padded = torch.zeros(width, dtype=torch.int64, device=dev)
        padded[: mine.numel()] = mine.to(dev)
        everyone = [torch.zeros_like(padded) for _ in range(self.world_size)]
        dist.all_gather(everyone, padded, group=self.process_group)
        self._collectives["allgather"] += 1
        # every rank checks every other rank, so all ranks raise together instead of one rank
        # erroring while its peers hang in the next collective
        for other, meta in enumerate(everyone):
            if other == self.rank:
                continue
            msg = _C.compare_param_metadata(padded.cpu(), meta.cpu(), other)
            if msg:
                raise RuntimeError(f"DDP expects the same model across all ranks, but rank {self.rank}: {msg}")

    def _sync_module_states(self, src: int = 0) -> None:
        ignored = set(getattr(self.module, "_ddp_params_and_buffers_to_ignore", ()) or ())
        params = [p.detach() for n, p in self.module.named_parameters() if n not in ignored]
        bufs = [b.detach() for n, b in self.module.named_buffers() if n not in ignored]
        self._collectives["broadcast"] += broadcast_coalesced(
            self._comm, params + bufs, self.broadcast_bucket_size, src=src)

    def _publish_grad_slots(self) -> None:
        """Expose each parameter's bucket slot so in-place wgrad kernels can target it."""
        for i, p in enumerate(self._params):
            p._grad_slot = self.reducer.grad_slot(i)
            p._ddp_index = i

    # -- forward ---------------------------------------------------------------------------
    def _pre_forward(self) -> None:
        if self._join_enabled and torch.is_grad_enabled() and self.require_backward_grad_sync:
            self._join_poll(1)              # tell ranks that already ran out of data to shadow this iteration
            self._join_iterations += 1
        if torch.is_grad_enabled() and self.require_backward_grad_sync:
            self.reducer.prepare_for_forward()
            if self.reducer.should_rebuild_buckets():
                self._rebuild_buckets()
        if self.broadcast_buffers and self.world_size > 1 and self._buffers_for_sync \
                and self.require_backward_grad_sync:
            if self._buffers_ahead and self._buffer_src == 0:
                self._buffers_ahead = False
                self._comm.wait_buffers_ahead()      # already broadcast under the previous backward
            else:
                self._sync_buffers()

    def _post_forward(self) -> None:
        """Issue the NEXT forward's buffer broadcast now, on the communication stream: rank 0's running statistics
        are final as soon as its forward is done, so K3 overlaps the backward pass instead of heading the next step
        (same values as the stock pre-forward broadcast, /root/reference/dpp.py:50 -> distributed.py _pre_forward)."""
        ahead = getattr(self._comm, "sync_buffers_ahead", None)
        if (ahead is None or self._join_enabled or self._buffer_src != 0 or not self.training
                or not torch.is_grad_enabled() or not self.will_sync_module_buffers()):
            return
        with torch.no_grad():
            self._collectives["broadcast"] += ahead(self.module, self._buffers_for_sync)
        self._buffers_ahead = True

    def load_state_dict(self, *args, **kwargs):
        self._buffers_ahead = False          # buffers changed behind the broadcast: the next forward syncs in line
        return super().load_state_dict(*args, **kwargs)

    def check_health(self, sync: bool = False) -> None:
        """Raise if a collective kernel on this rank gave up waiting for a peer (dead or stalled rank).  Reads one
        word of pinned host memory: cheap enough to call after every step (``sync=True`` first drains the GPU)."""
        comm = self._comm
        while comm is not None:
            check = getattr(comm, "check", None)
            if check is not None:
                check(sync=sync)
            comm = getattr(comm, "inner", None)
        if sync and self.device.type == "cuda":
            # the tcgen05 kernels' own wait-timeout word (pipeline barriers, split-K tickets): device memory, so it is
            # read only at the synchronising call sites (log interval with sync, checkpoints, clean-up)
            from ..ops import native

            w = native.tc_error_word(self.device)
            if w:
                raise RuntimeError(f"a tensor-core kernel on rank {self.rank} gave up waiting inside its pipeline "
                                   f"(error word {w:#x}): results of this step are not trustworthy")

    def _rebuild_buckets(self) -> None:
        order = torch.tensor(self.reducer.rebuild_order(), dtype=torch.int64)
        if self.world_size > 1:
            dev = self.device if dist.get_backend(self.process_group) == "nccl" else torch.device("cpu")
            order = order.to(dev)
            count = torch.tensor([order.numel()], dtype=torch.int64, device=dev)
            dist.broadcast(count, src=0, group=self.process_group)     # K5: sizes ...
            dist.broadcast(order, src=0, group=self.process_group)     # ... and rank 0's order
            self._collectives["broadcast"] += 2
        self.reducer.rebuild_buckets(order.cpu().tolist())
        if self.device.type == "cuda" and self.world_size > 1:
            # iteration 0 has shown every convolution shape to the tuner: agree on ONE table of kernel variants so
            # that all ranks run the same code from here on (and persist it for the next run)
            from ..ops import native

            native.sync_choices(self.process_group)
            native.save_choices()
        self._rebuilt_bucket_sizes = list(self.reducer.bucket_sizes())
        self._publish_grad_slots()

    def _sync_buffers(self) -> None:
        with torch.no_grad():
            sync = getattr(self._comm, "sync_buffers", None)
            if sync is not None and self._buffer_src == 0:
                self._collectives["broadcast"] += sync(self.module, self._buffers_for_sync)
            else:
                self._collectives["broadcast"] += broadcast_coalesced(
                    self._comm, self._buffers_for_sync, self.broadcast_bucket_size, src=self._buffer_src)

    # -- device-timed logger ---------------------------------------------------------------
    FWD_START, BWD_START, BWD_END, STEP_END = 0, 1, 2, 3

    def enable_device_timing(self, on: bool = True) -> None:
        """Stamp %globaltimer on the training stream at forward start / backward start / backward end (and step end,
        by ``TrainStep``).  The marks are 1-thread kernels, so they are captured in a step graph and keep working
        under replay -- enable BEFORE the step is captured.  The collective kernels stamp their own start / end for
        free; together they give device-timed forward / backward / communication / exposed-communication figures
        (``device_timing()``, ``--print-ddp-stats``), replacing host clocks that measure nothing under a CUDA graph."""
        self._timing = bool(on) and self.device.type == "cuda"
        if self._timing and self._marks is None:
            self._marks = torch.zeros(8, dtype=torch.int64, device=self.device)
            self._step_timer = _C.StepTimer(0, 1)

    def _mark(self, slot: int) -> None:
        if self._timing:
            from ..ops import native

            native.stamp(self._marks.data_ptr() + 8 * slot)

    def device_timing(self) -> dict:
        """Figures of the LAST step, from device time stamps (synchronises): forward_us, backward_us, comm_span_us,
        exposed_comm_us (= last bucket reduction end - backward compute end, clamped at 0), exposed_comm_pct of the
        step, and per-launch-unit duration / bus GB/s."""
        comm = self._comm
        while getattr(comm, "inner", None) is not None:
            comm = comm.inner
        stamps, nbytes, bcast = comm.raw_stamps() if hasattr(comm, "raw_stamps") else ([], [], (0, 0))
        marks = [0, 0, 0, 0]
        if self._marks is not None:
            torch.cuda.synchronize(self.device)
            marks = self._marks.cpu().tolist()[:4]
        # the arithmetic (which reductions belong to this step, span, busy, exposed) lives in the C++ timer
        out = dict(_C.StepTimer.summarize(marks, stamps, nbytes, self.world_size))
        if bcast[0] and bcast[1] >= bcast[0]:
            out["buffer_broadcast_us"] = (bcast[1] - bcast[0]) / 1e3
        if self._step_timer is not None:
            self._step_timer.record(marks, stamps, nbytes)
        return out

    def forward(self, *inputs, **kwargs):
        with torch.autograd.profiler.record_function("DistributedDataParallel.forward"):
            self._pre_forward()
            self._mark(self.FWD_START)
            out = self.module(*inputs, **kwargs)
            self._post_forward()
            self._mark(self.BWD_START)
            if torch.is_grad_enabled():
                for p in self._params:
                    p._slot_claimed = False      # one in-place writer per parameter per backward (ops/native.py)
            if torch.is_grad_enabled() and self.require_backward_grad_sync:
                outs = [t for t in _flatten(out) if isinstance(t, torch.Tensor)] if self.find_unused_parameters else []
                self.reducer.prepare_for_backward(outs)
            return out

    # -- public DDP surface ------------------------------------------------------------------
    @contextlib.contextmanager
    def no_sync(self):
        """Skip gradient synchronisation inside the context; gradients accumulate locally and
        are reduced by the first backward after leaving it (distributed.py:1474-1500)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        self.reducer.set_require_sync(False)
        try:
            yield
        finally:
            self.require_backward_grad_sync = old
            self.reducer.set_require_sync(old)

    # -- uneven inputs ----------------------------------------------------------------------
    def _control_tensor(self, values) -> torch.Tensor:
        dev = self.device if dist.get_backend(self.process_group) == "nccl" else torch.device("cpu")
        return torch.tensor(values, dtype=torch.int64, device=dev)

    def _join_poll(self, active: int) -> int:
        """One tiny all-reduce per iteration while ``join()`` is active: a one-hot vector of the ranks that still
        have data.  Returns the highest such rank (-1: none) -- that rank is the source of this iteration's
        buffer broadcast, as in stock DDP -- and sets the gradient divisor for the iteration."""
        t = self._control_tensor([1 if (active and r == self.rank) else 0 for r in range(self.world_size)])
        dist.all_reduce(t, group=self.process_group)
        self._collectives["allreduce_control"] += 1
        flags = t.tolist()
        n_active = int(sum(flags))
        top = max((r for r, f in enumerate(flags) if f), default=-1)
        self._buffer_src = max(top, 0)
        if not self._join_divide_by_world and n_active > 0:
            self._set_divisor(n_active)
        if self._join_throw and 0 < n_active < self.world_size:
            # every rank sees the same vector in the same iteration, so every rank raises here together
            self._join_enabled = False
            self._set_divisor(self.world_size)
            raise RuntimeError(f"Detected at least one rank that exhausted inputs. Throwing across all ranks "
                               f"(rank {self.rank}; {n_active} of {self.world_size} ranks still had data).")
        return top

    def _set_divisor(self, value: int) -> None:
        comm = self._comm
        while comm is not None:                      # a comm hook wraps the data plane it was registered over
            comm.divisor = value
            comm = getattr(comm, "inner", None)

    def _shadow_iteration(self) -> None:
        """What a rank without data does while others train: the same collectives, zero gradients."""
        self.reducer.prepare_for_forward()
        if self.reducer.should_rebuild_buckets():
            self._rebuild_buckets()
        if self.broadcast_buffers and self._buffers_for_sync:
            self._sync_buffers()
        self.reducer.shadow_backward()

    @contextlib.contextmanager
    def join(self, divide_by_initial_world_size: bool = True, enable: bool = True,
             throw_on_early_termination: bool = False):
        """Train with a different number of batches per rank (``DistributedDataParallel.join``,
        distributed.py:1798).  A rank that leaves the context first keeps matching the other ranks'
        collectives with zero gradients; gradients are divided by the full world size, or -- with
        ``divide_by_initial_world_size=False`` -- by the number of ranks that still have data (the factor is
        an argument of the reduction kernels, set per iteration).  When the last rank leaves, its parameters
        and buffers are broadcast so every replica ends identical.  With ``throw_on_early_termination=True`` all
        ranks raise ``RuntimeError`` in the iteration in which the first rank runs out instead.  Steps run eagerly while the context is
        active (a captured CUDA graph cannot carry the per-iteration control collective)."""
        if not enable or self.world_size == 1:
            yield
            return
        self._join_enabled = True
        self._join_iterations = 0
        self._join_divide_by_world = bool(divide_by_initial_world_size)
        self._join_throw = bool(throw_on_early_termination)
        try:
            yield
        except BaseException:
            self._join_enabled = False
            self._set_divisor(self.world_size)
            raise
        while self._join_poll(0) >= 0:
            self._shadow_iteration()
        self._join_enabled = False
        self._buffer_src = 0
        self._set_divisor(self.world_size)
        # the replica that trained longest (highest rank among ties) is authoritative
        key = self._control_tensor([self._join_iterations * self.world_size + self.rank])
        dist.all_reduce(key, op=dist.ReduceOp.MAX, group=self.process_group)
        self._collectives["allreduce_control"] += 1
        self._sync_module_states(src=int(key.item()) % self.world_size)

    def register_comm_hook(self, state, hook: Callable) -> None:
        """``hook(state, bucket: GradBucket) -> Future[Tensor]``; the hook owns the averaging."""
        if not callable(hook):
            raise TypeError("Communication hook must be callable.")
        if self._comm_hook_name != "builtin":
            raise RuntimeError("register_comm_hook can only be called once.")
        self._comm_hook_name = getattr(hook, "__qualname__", repr(hook))
        wrapped = HookComm(self.process_group, state, hook, self._bucket_info, self._comm)
        self._comm = wrapped
        self.reducer.set_comm(wrapped)

    def _bucket_info(self, index: int):
        members = self.reducer.bucket_params()[index]
        return [self.reducer.grad_slot(i) for i in members], [self._params[i] for i in members]

    def _get_ddp_logging_data(self) -> dict:
        d = dict(self.reducer.logging_data())
        backend = dist.get_backend(self.process_group) if dist.is_initialized() else "none"
        d.update(
            module_name=type(self.module).__name__,
            backend_name=backend,
            comm_backend=self._comm.name,
            comm_hook=self._comm_hook_name,
            device_ids=",".join(map(str, self.device_ids or [])),
            output_device=-1 if self.output_device is None else self.output_device,
            broadcast_buffers=self.broadcast_buffers,
            rank=self.rank,
            initial_bucket_sizes=self._initial_bucket_sizes,
            rebuilt_bucket_sizes=self._rebuilt_bucket_sizes,
            launch_unit_sizes=list(self.reducer.launch_unit_sizes()),
            collectives=dict(self._collectives),
            bucket_launches=self._comm.launches,
        )
        stats = getattr(self._comm, "stats", None)
        if stats is not None:
            d["comm_stats"] = stats()
        if self._step_timer is not None and self._step_timer.samples() > 0:
            d.update(self._step_timer.averages())      # device-timed averages (forward / backward / exposed comm)
        return d

    @property
    def comm(self) -> CommBase:
        return self._comm

    def parameters_in_bucket_order(self) -> List[torch.Tensor]:
        return [self._params[i] for b in self.reducer.bucket_params() for i in b]

    def __del__(self):
        red = self.__dict__.get("reducer")
        if red is not None:
            try:
                red.detach_hooks()
            except Exception:
                pass


def _device_index(d) -> int:
    if isinstance(d, torch.device):
        return d.index if d.index is not None else torch.cuda.current_device()
    return int(d)


def _flatten(obj):
    if isinstance(obj, (list, tuple)):
        for o in obj:
            yield from _flatten(o)
    elif isinstance(obj, dict):
        for o in obj.values():
            yield from _flatten(o)
    else:
        yield obj
