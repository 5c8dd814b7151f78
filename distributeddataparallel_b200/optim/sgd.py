"""Fused SGD (momentum / dampening / weight decay / nesterov / maximize)."""
from __future__ import annotations

import torch

from .. import ops


class SGD(torch.optim.SGD):
    """``optim.SGD(model.parameters(), lr=0.01)`` (/root/reference/dpp.py:41) with a fused step.

    ``state_dict()`` is the stock one; momentum buffers are stored under
    ``state[p]["momentum_buffer"]`` exactly like torch (T/optim/sgd.py:252-320).
    """

    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0,
                 nesterov=False, *, maximize: bool = False, native: bool = True, **kw):
        kw.pop("foreach", None)
        kw.pop("fused", None)
        super().__init__(params, lr=lr, momentum=momentum, dampening=dampening,
                         weight_decay=weight_decay, nesterov=nesterov, maximize=maximize)
        self._native = native
        self._tables = {}

    def supports_subset_steps(self) -> bool:
        """One param group on CUDA with the native kernels: the update can be applied bucket by bucket."""
        return bool(self._native and len(self.param_groups) == 1 and ops.native_enabled()
                    and all(p.is_cuda for p in self.param_groups[0]["params"]))

    @torch.no_grad()
    def step_subset(self, key, params) -> None:
        """The SGD update for ``params`` only (those that have a gradient), on the current stream.  Used by
        ``engine.TrainStep`` to update a gradient bucket's parameters as soon as that bucket is reduced, while the
        backward pass is still producing the other buckets; the remaining parameters are updated by ``step()`` with
        ``skip=`` set to the ones already done."""
        self._update(("subset", key), self.param_groups[0], [p for p in params if p.grad is not None], join=False)

    @torch.no_grad()
    def step(self, closure=None, skip=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not self._native:
            return super().step()
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if p.grad is not None and (skip is None or id(p) not in skip)]
            if not ps:
                continue
            self._update(gi, group, ps)
        return loss

    def _update(self, gi, group, ps, join: bool = True) -> None:
        if not ps:
            return
        if not (ps[0].is_cuda and ops.native_enabled()):
            self._reference_step(group, ps)
            return
        from ..ops import native

        if join:
            native.join_side_streams(ps[0].device)   # weight gradients may still be in flight on a side stream
        mom = group["momentum"]
        bufs = None
        first = False
        if mom != 0:
            fresh = [p for p in ps if self.state[p].get("momentum_buffer") is None]
            # torch initialises a momentum buffer with the first gradient (no dampening).  The fused kernel
            # does that for the whole group through one flag, so it is used only when EVERY buffer is new;
            # a parameter that joins later gets its buffer initialised here, on its own.
            first = len(fresh) == len(ps)
            for p in fresh:
                if first:
                    buf = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.preserve_format)
                else:
                    g = -p.grad if group["maximize"] else p.grad
                    if group["weight_decay"] != 0:
                        g = g.add(p, alpha=group["weight_decay"])
                    buf = g.detach().to(torch.float32).clone(memory_format=torch.preserve_format)
                    buf.div_(mom)        # the kernel's buf = mom * buf + (1 - dampening) * g is then undone below
                    buf.sub_(g.to(torch.float32), alpha=(1 - group["dampening"]) / mom)
                self.state[p]["momentum_buffer"] = buf
            bufs = [self.state[p]["momentum_buffer"] for p in ps]
        native.fused_sgd(self._tables, gi, ps, [p.grad for p in ps], bufs, lr=group["lr"],
                         momentum=mom, dampening=group["dampening"], weight_decay=group["weight_decay"],
                         nesterov=group["nesterov"], maximize=group["maximize"], first_step=first)

    def sync_hyper(self) -> None:
        """Push changed hyper-parameters (a scheduler moved ``lr``) to the device-side table a captured step reads."""
        for key, tab in self._tables.items():
            group = self.param_groups[key if isinstance(key, int) else 0]
            tab[1].set_hyper((group["lr"], group["momentum"], group["dampening"], group["weight_decay"],
                              float(group["nesterov"]), float(group["maximize"]), 0.0, 1.0))

    def load_state_dict(self, state_dict):
        """Stock loader, then momentum buffers back to fp32: ``torch.optim.Optimizer.load_state_dict`` casts
        floating-point state to the parameter's dtype, and the fused kernels keep fp32 state for bf16 params."""
        super().load_state_dict(state_dict)
        for st in self.state.values():
            buf = st.get("momentum_buffer")
            if torch.is_tensor(buf) and buf.is_cuda and buf.dtype != torch.float32:
                st["momentum_buffer"] = buf.to(torch.float32)
        self._tables.clear()

    @staticmethod
    def _plain(group, p, g, state):
        if group["maximize"]:
            g = -g
        if group["weight_decay"] != 0:
            g = g.add(p, alpha=group["weight_decay"])
        mom = group["momentum"]
        if mom != 0:
            buf = state.get("momentum_buffer")
            if buf is None:
                buf = state["momentum_buffer"] = g.clone()
            else:
                buf.mul_(mom).add_(g, alpha=1 - group["dampening"])
            g = g.add(buf, alpha=mom) if group["nesterov"] else buf
        p.add_(g, alpha=-group["lr"])

    def _reference_step(self, group, ps):
        for p in ps:
            self._plain(group, p, p.grad, self.state[p])
