"""Fused Adam / AdamW."""
from __future__ import annotations

import math

import torch

from .. import ops


class Adam(torch.optim.Adam):
    """Adam with the torch state layout (``step``, ``exp_avg``, ``exp_avg_sq``) and one fused
    multi-tensor launch per param group on CUDA.  ``decoupled=True`` gives AdamW."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
                 *, decoupled: bool = False, maximize: bool = False, native: bool = True, **kw):
        kw.pop("foreach", None)
        kw.pop("fused", None)
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, maximize=maximize)
        self._native = native
        self._decoupled = decoupled
        self._tables = {}

    def sync_hyper(self) -> None:
        """Push changed hyper-parameters to the device-side table a captured step reads."""
        for gi, group in enumerate(self.param_groups):
            tab = self._tables.get(gi)
            if tab is not None:
                b1, b2 = group["betas"]
                tab[1].set_hyper((group["lr"], b1, b2, group["eps"], group["weight_decay"], float(self._decoupled),
                                  float(group["maximize"]), 1.0))

    def load_state_dict(self, state_dict):
        """Stock loader, then moments back to fp32 (the stock loader casts state to the parameter dtype; the fused
        kernel keeps fp32 moments for bf16 parameters)."""
        super().load_state_dict(state_dict)
        for st in self.state.values():
            for k in ("exp_avg", "exp_avg_sq"):
                v = st.get(k)
                if torch.is_tensor(v) and v.is_cuda and v.dtype != torch.float32:
                    st[k] = v.to(torch.float32)
            if torch.is_tensor(st.get("step")) and st["step"].dtype != torch.float32:
                st["step"] = st["step"].to(torch.float32)
        self._tables.clear()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            for p in ps:
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.zeros((), dtype=torch.float32, device=p.device if p.is_cuda else "cpu")
                    sdt = torch.float32 if p.is_cuda else p.dtype      # fused kernels keep fp32 state
                    st["exp_avg"] = torch.zeros_like(p, dtype=sdt, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=sdt, memory_format=torch.preserve_format)
            b1, b2 = group["betas"]
            if self._native and ps[0].is_cuda and ops.native_enabled():
                from ..ops import native

                native.join_side_streams(ps[0].device)
                native.fused_adam(self._tables, gi, ps, [p.grad for p in ps],
                                  [self.state[p]["exp_avg"] for p in ps],
                                  [self.state[p]["exp_avg_sq"] for p in ps],
                                  [self.state[p]["step"] for p in ps],
                                  lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"],
                                  weight_decay=group["weight_decay"], decoupled=self._decoupled,
                                  maximize=group["maximize"])
                continue
            for p in ps:
                st = self.state[p]
                g = -p.grad if group["maximize"] else p.grad
                st["step"] += 1
                t = float(st["step"])
                if group["weight_decay"] != 0:
                    if self._decoupled:
                        p.mul_(1 - group["lr"] * group["weight_decay"])
                    else:
                        g = g.add(p, alpha=group["weight_decay"])
                st["exp_avg"].lerp_(g, 1 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1 - b2)
                bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
                denom = (st["exp_avg_sq"].sqrt() / math.sqrt(bc2)).add_(group["eps"])
                p.addcdiv_(st["exp_avg"], denom, value=-group["lr"] / bc1)
        return loss
