"""One training step, eager or replayed from a CUDA graph.

The reference's step is seven Python statements (/root/reference/dpp.py:48-53) that expand
into ~300 kernel launches for 7 GFLOP of math ("the stock stack is
launch-/latency-bound on B200").  ``TrainStep`` keeps the same semantics

    optimizer.zero_grad(); out = model(x); loss = criterion(out, y); loss.backward(); optimizer.step()

but, once bucket layout and shapes are final (after the one-time bucket rebuild of
iteration 1), captures the *whole* step -- forward, backward, the reducer's overlapped
bucket all-reduces on the side stream, and the fused optimizer update -- into one CUDA
graph and replays it.  Hooks run at capture time only; what is replayed is the kernel
DAG they produced, including the fork/join onto the communication stream.

Shapes: the steady-state batch is captured first; a second shape -- the ragged last batch of
an epoch (dpp.py:35 keeps ``drop_last=False``: 16 / 8 / 20 / 10 samples at 1 / 2 / 4 / 8 ranks,
SURVEY §3.4), identical on every rank -- runs eagerly the first time it is seen (the
convolution tuner needs one eager pass per new geometry) and is captured the second time, up
to ``max_graphs`` graphs (``B200DDP_MAX_GRAPHS``, default 2).

Optimizer in backward: with this framework's fused SGD (one param group) and the P2P / single-process data plane,
the update of a gradient bucket's parameters is enqueued right behind that bucket's reduction on the communication
stream (``comm.after_reduce``), so it runs under the rest of the backward pass; only the few KB of the last launch
unit are updated after backward ends.  Same arithmetic as ``optimizer.step()``, earlier in time (nothing in the
backward pass reads a parameter after its gradient is complete).  ``overlap_optimizer=False`` switches it off.

After every step the DDP wrapper's health word (pinned host memory, written by a collective
kernel whose peer never arrived) is read; a set word raises on this rank.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Tuple

import torch
import torch.nn as nn


class TrainStep:
    def __init__(self, model: nn.Module, criterion: Callable, optimizer: torch.optim.Optimizer,
                 cuda_graph: bool = True, warmup_steps: int = 3, overlap_optimizer: bool = True) -> None:
        self.model = model
        self.criterion = criterion
        self.optimizer = optimizer
        self.warmup_steps = max(3, int(warmup_steps))
        self.steps = 0
        self.replays = 0
        self.eager_steps = 0
        self._graphs: Dict[Tuple, dict] = {}
        self._seen: Dict[Tuple, int] = {}
        dev = next(model.parameters()).device
        self.use_graph = bool(cuda_graph) and dev.type == "cuda"
        self._ddp = model if hasattr(model, "reducer") else None
        import os

        self.max_graphs = int(os.environ.get("B200DDP_MAX_GRAPHS", "2"))
        self._updated = set()          # ids of the parameters already updated by per-bucket optimizer launches
        self._overlap_opt = False
        if (overlap_optimizer and os.environ.get("B200DDP_OVERLAP_OPT", "1") != "0" and self._ddp is not None
                and dev.type == "cuda" and getattr(optimizer, "supports_subset_steps", lambda: False)()
                and getattr(self._ddp.comm, "name", "") in ("p2p", "none")):
            import weakref

            me = weakref.ref(self)
            self._opt_ids = {id(p) for g in optimizer.param_groups for p in g["params"]}
            self._ddp.comm.after_reduce = lambda index: me() is not None and me()._after_reduce(index)
            self._overlap_opt = True
        self._nvtx = dev.type == "cuda" and os.environ.get("B200DDP_NVTX", "0") != "0"

    def _after_reduce(self, index: int) -> None:
        """Runs on the stream that carries launch unit ``index``'s reduction, right behind it."""
        ddp = self._ddp
        if ddp._join_enabled or not ddp.training:
            return                   # uneven-input shadow iterations / evaluation: the ordinary step() decides
        ps = [ddp._params[i] for i in ddp.reducer.bucket_params()[index]]
        ps = [p for p in ps if id(p) in self._opt_ids and p.grad is not None]
        if ps:
            self.optimizer.step_subset(index, ps)
            self._updated.update(id(p) for p in ps)

    def _optimizer_step(self) -> None:
        if self._overlap_opt and self._updated:
            done, self._updated = self._updated, set()
            self.optimizer.step(skip=done)          # whatever no bucket launch covered (nothing in steady state)
        else:
            self.optimizer.step()

    def _forward_loss(self, data, target):
        inner = getattr(self.model, "module", self.model)
        if getattr(inner, "computes_loss", False):          # e.g. GPT-2: fused LM head + cross-entropy
            return self.model(data, target)
        return self.criterion(self.model(data), target)

    # -- eager ---------------------------------------------------------------------------
    def _eager(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        nvtx = self._nvtx
        self._updated = set()
        self.optimizer.zero_grad()
        if nvtx:
            torch.cuda.nvtx.range_push("b200ddp.forward")
        loss = self._forward_loss(data, target)
        if nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("b200ddp.backward+reduce")
        loss.backward()
        if nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("b200ddp.optimizer")
        self._optimizer_step()
        if self._ddp is not None:
            self._ddp._mark(self._ddp.STEP_END)
        if nvtx:
            torch.cuda.nvtx.range_pop()
        self.eager_steps += 1
        return loss.detach()

    # -- graph ---------------------------------------------------------------------------
    def _ready_for_capture(self) -> bool:
        if self.steps < self.warmup_steps:
            return False
        if self._ddp is not None and not self._ddp.find_unused_parameters:
            return bool(self._ddp.reducer.has_rebuilt())
        return True

    def _capture(self, data: torch.Tensor, target: torch.Tensor) -> dict:
        static_x = data.clone(memory_format=torch.preserve_format)
        static_y = target.clone()
        self._updated = set()
        self.optimizer.zero_grad(set_to_none=True)
        from .ops import native

        native.join_side_streams()
        native.save_choices()
        torch.cuda.synchronize()
        comm = getattr(self._ddp, "comm", None)
        if comm is not None and getattr(comm, "_ahead_event", None) is not None:
            comm._ahead_event = None       # everything issued so far is complete: nothing to wait for inside the capture

        launches0 = native.launch_count()
        graph = torch.cuda.CUDAGraph()
        cap_stream = torch.cuda.Stream()
        cap_stream.wait_stream(torch.cuda.current_stream())
        if self._ddp is not None:
            # autograd pins a gradient accumulator to the stream it was created on: recreate them
            # under the capture stream (and put them back on the caller's stream afterwards)
            with torch.cuda.stream(cap_stream):
                self._ddp.reducer.attach_hooks()
        try:
            # thread_local: the data loader's staging thread keeps issuing copies while the step is captured
            with torch.cuda.graph(graph, stream=cap_stream, capture_error_mode="thread_local"):
                loss = self._forward_loss(static_x, static_y)
                loss.backward()
                self._optimizer_step()
                if self._ddp is not None:
                    self._ddp._mark(self._ddp.STEP_END)
                # drop the autograd graph now: it keeps the capture-stream accumulator nodes alive
                loss = loss.detach()
        finally:
            if self._ddp is not None:
                self._ddp.reducer.attach_hooks()
            if comm is not None and getattr(comm, "_ahead_event", None) is not None:
                comm._ahead_event = None   # recorded inside the capture: the graph's own join orders later work
        native.set_launches_per_capture(native.launch_count() - launches0)
        entry = {"graph": graph, "x": static_x, "y": static_y, "loss": loss.detach()}
        return entry

    def __call__(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        self.steps += 1
        if not self.use_graph or not self.model.training or (self._ddp is not None and self._ddp._join_enabled):
            loss = self._eager(data, target)         # join(): every iteration carries a control collective
            self._health()
            return loss
        key = (tuple(data.shape), data.dtype, tuple(target.shape))
        entry = self._graphs.get(key)
        if entry is None:
            seen = self._seen.get(key, 0)
            self._seen[key] = seen + 1
            # a shape other than the first captured one (ragged tail) gets one eager pass first
            first_pass = bool(self._graphs) and seen == 0
            if not self._ready_for_capture() or len(self._graphs) >= self.max_graphs or first_pass:
                loss = self._eager(data, target)
                self._health()
                return loss
            entry = self._graphs[key] = self._capture(data, target)
            # the capture itself does not execute the step: fall through to a replay
        sync_hyper = getattr(self.optimizer, "sync_hyper", None)
        if sync_hyper is not None:
            sync_hyper()                 # the captured update reads lr & co. from device memory: refresh if changed
        entry["x"].copy_(data, non_blocking=True)
        entry["y"].copy_(target, non_blocking=True)
        entry["graph"].replay()
        self.replays += 1
        if self._ddp is not None:
            self._ddp.reducer.note_graph_replay()
            self._health()
        return entry["loss"]

    def _health(self) -> None:
        if self._ddp is not None:
            self._ddp.check_health(sync=False)

    def load_inputs_and_replay(self, entry_key=None) -> torch.Tensor:
        """Replay on whatever is already in the static input buffers (bench helper)."""
        entry = next(iter(self._graphs.values())) if entry_key is None else self._graphs[entry_key]
        entry["graph"].replay()
        self.replays += 1
        return entry["loss"]

    @property
    def captured(self) -> bool:
        return bool(self._graphs)

    def static_inputs(self):
        entry = next(iter(self._graphs.values()))
        return entry["x"], entry["y"]
