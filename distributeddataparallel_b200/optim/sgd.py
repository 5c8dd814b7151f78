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

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not self._native:
            return super().step()
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            if not (ps[0].is_cuda and ops.native_enabled()):
                self._reference_step(group, ps)
                continue
            from ..ops import native

            mom = group["momentum"]
            bufs = None
            first = False
            if mom != 0:
                bufs = []
                for p in ps:
                    st = self.state[p]
                    if st.get("momentum_buffer") is None:
                        # torch initialises the buffer with the first gradient (no dampening)
                        st["momentum_buffer"] = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.preserve_format)
                        first = True
                    bufs.append(st["momentum_buffer"])
            native.fused_sgd(self._tables, gi, ps, [p.grad for p in ps], bufs, lr=group["lr"],
                             momentum=mom, dampening=group["dampening"], weight_decay=group["weight_decay"],
                             nesterov=group["nesterov"], maximize=group["maximize"], first_step=first)
        return loss

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
