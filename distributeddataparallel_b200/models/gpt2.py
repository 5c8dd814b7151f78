"""GPT-2 (small / medium / tiny test size) for the driver-specified DDP configuration
(BASELINE.json config 4: "GPT-2 medium DDP bf16 with 25 MB gradient buckets").  Not part of the
reference script (dpp.py trains a CNN); sized per SURVEY.md §6.3: 24 layers, d_model 1024,
16 heads, vocabulary 50 257, context 1024, ~354.8 M parameters with the tied LM head.

On CUDA every op is ours: tcgen05 GEMMs for the QKV / projection / MLP / LM-head products and the
batched QK^T / PV of attention, hand-written LayerNorm, GELU, causal softmax, embedding and
cross-entropy kernels (ops/native_tf.py).  On CPU the same module runs on PyTorch ops (tests).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops


@dataclass
class GPT2Config:
    n_layer: int = 24
    n_head: int = 16
    d_model: int = 1024
    vocab_size: int = 50257
    block_size: int = 1024


CONFIGS = {
    "gpt2": GPT2Config(12, 12, 768),
    "gpt2-small": GPT2Config(12, 12, 768),
    "gpt2-medium": GPT2Config(24, 16, 1024),
    "gpt2-tiny": GPT2Config(2, 4, 128, 512, 64),
}


def _native(t: torch.Tensor) -> bool:
    return ops.use_native(t)


class Linear(nn.Linear):
    def forward(self, x):
        if _native(x):
            from ..ops import native_tf

            return native_tf.linear_tc(x, self.weight, self.bias)
        return F.linear(x, self.weight, self.bias)


class LayerNorm(nn.LayerNorm):
    def forward(self, x):
        if _native(x):
            from ..ops import native_tf

            return native_tf.layer_norm(x, self.weight, self.bias, self.eps)
        return F.layer_norm(x, self.normalized_shape, self.weight, self.bias, self.eps)


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config) -> None:
        super().__init__()
        d = cfg.d_model
        self.n_head = cfg.n_head
        self.ln_1 = LayerNorm(d)
        self.c_attn = Linear(d, 3 * d)
        self.c_proj = Linear(d, d)
        self.ln_2 = LayerNorm(d)
        self.c_fc = Linear(d, 4 * d)
        self.c_proj2 = Linear(4 * d, d)

    def attention(self, qkv: torch.Tensor) -> torch.Tensor:
        if _native(qkv):
            from ..ops import native_tf

            return native_tf.causal_self_attention(qkv, self.n_head)
        b, t, three_d = qkv.shape
        d = three_d // 3
        q, k, v = qkv.view(b, t, 3, self.n_head, d // self.n_head).permute(2, 0, 3, 1, 4)
        o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return o.transpose(1, 2).reshape(b, t, d)

    def forward(self, x):
        if _native(x) and self.c_fc.bias is not None:
            from ..ops import native_tf

            # residual adds, GELU and its derivative ride in GEMM epilogues; the residual branches' gradients
            # are added inside the LayerNorm backward kernels
            h, xr = native_tf.layer_norm_res(x, self.ln_1.weight, self.ln_1.bias, self.ln_1.eps)
            a = self.attention(self.c_attn(h))
            x = native_tf.linear_tc(a, self.c_proj.weight, self.c_proj.bias, residual=xr)
            h, xr = native_tf.layer_norm_res(x, self.ln_2.weight, self.ln_2.bias, self.ln_2.eps)
            return native_tf.mlp(h, self.c_fc.weight, self.c_fc.bias, self.c_proj2.weight, self.c_proj2.bias, residual=xr)
        x = x + self.c_proj(self.attention(self.c_attn(self.ln_1(x))))
        h = F.gelu(self.c_fc(self.ln_2(x)), approximate="tanh")
        return x + self.c_proj2(h)


class GPT2(nn.Module):
    computes_loss = True      # forward(idx, targets) returns the loss (fused LM head + cross-entropy)

    def __init__(self, cfg: GPT2Config) -> None:
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.vocab_size, cfg.d_model)
        self.wpe = nn.Embedding(cfg.block_size, cfg.d_model)
        self.h = nn.ModuleList(Block(cfg) for _ in range(cfg.n_layer))
        self.ln_f = LayerNorm(cfg.d_model)
        self.apply(self._init)
        for blk in self.h:                                   # GPT-2 residual-projection scaling
            for p in (blk.c_proj.weight, blk.c_proj2.weight):
                nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                nn.init.zeros_(m.bias)

    def forward(self, idx: torch.Tensor, targets: torch.Tensor | None = None):
        b, t = idx.shape
        if _native(self.wte.weight):
            from ..ops import native_tf

            x = native_tf.embedding(idx, self.wte.weight, self.wpe.weight)
        else:
            x = self.wte(idx) + self.wpe(torch.arange(t, device=idx.device))
        for blk in self.h:
            x = blk(x)
        x = self.ln_f(x)
        if targets is None:
            return F.linear(x, self.wte.weight)
        if _native(x):
            from ..ops import native_tf

            return native_tf.lm_head_loss(x, self.wte.weight, targets)
        return F.cross_entropy(F.linear(x, self.wte.weight).float().view(b * t, -1), targets.reshape(-1))

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())


def build_gpt2(name: str = "gpt2-medium", seq_len: int = 1024) -> GPT2:
    cfg = CONFIGS[name]
    if seq_len and seq_len != cfg.block_size:
        cfg = GPT2Config(cfg.n_layer, cfg.n_head, cfg.d_model, cfg.vocab_size, seq_len)
    return GPT2(cfg)
