"""Functional op layer.

Every op has two implementations:

* the **native** one -- hand-written sm_100a kernels from ``csrc/kernels`` reached through
  the C ABI declared in ``include/b200ddp.h`` (``ops/native.py`` holds the autograd
  Functions) -- used whenever the tensors live on a CUDA device;
* a **plain PyTorch** one used on CPU (gloo plumbing tests) and as the fp32 oracle in the
  numerics tests.

On a GPU the native path is mandatory: if the kernel library cannot be loaded the ops
raise instead of silently falling back (set ``B200DDP_NATIVE=0`` to opt out explicitly,
e.g. to time the library ops for comparison).

These replace the library kernels the reference reaches through torchvision's ResNet and
``nn.CrossEntropyLoss`` (/root/reference/dpp.py:11-18, :40, :50-52; inventory in
SURVEY.md §2.5(b) rows K7-K17).
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn.functional as F

_FORCE_OFF = os.environ.get("B200DDP_NATIVE", "1") == "0"
_native_enabled = True


def set_native(enabled: bool) -> None:
    global _native_enabled
    _native_enabled = bool(enabled)


def native_enabled() -> bool:
    return _native_enabled and not _FORCE_OFF


def use_native(t: torch.Tensor) -> bool:
    return t.is_cuda and native_enabled()


def _nat():
    from . import native

    return native


# ------------------------------------------------------------------------------------
# input pipeline
# ------------------------------------------------------------------------------------
def preprocess_u8(x_u8: torch.Tensor) -> torch.Tensor:
    """uint8 NHWC -> normalised fp32, logical NCHW / physical NHWC (dpp.py:32 transform)."""
    if use_native(x_u8):
        return _nat().preprocess_u8(x_u8)
    from ..data import normalize_u8

    return normalize_u8(x_u8).permute(0, 3, 1, 2)


# ------------------------------------------------------------------------------------
# model math
# ------------------------------------------------------------------------------------
def conv2d(x: torch.Tensor, weight: torch.Tensor, stride: int, padding: int,
           weight_param: Optional[torch.Tensor] = None) -> torch.Tensor:
    if use_native(x):
        return _nat().conv2d(x, weight, stride, padding)
    return F.conv2d(x, weight, None, stride, padding)


def batch_norm_act(x, weight, bias, running_mean, running_var, num_batches_tracked,
                   residual=None, relu=False, training=True, momentum=0.1, eps=1e-5, ws=None):
    """BatchNorm2d (+ residual add) (+ ReLU) as one op; updates running stats in training."""
    if use_native(x):
        return _nat().batch_norm_act(x, weight, bias, running_mean, running_var, num_batches_tracked,
                                     residual, relu, training, momentum, eps, ws)
    if training and num_batches_tracked is not None:
        num_batches_tracked.add_(1)
    y = F.batch_norm(x, running_mean, running_var, weight, bias, training, momentum, eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


def max_pool2d(x, kernel_size=3, stride=2, padding=1):
    if use_native(x) and (kernel_size, stride, padding) == (3, 2, 1):
        return _nat().max_pool_3x3s2(x)
    return F.max_pool2d(x, kernel_size, stride, padding)


def global_avg_pool(x):
    """AdaptiveAvgPool2d((1,1)) + flatten -> [N, C]."""
    if use_native(x):
        return _nat().global_avg_pool(x)
    return x.mean(dim=(2, 3))


def linear(x, weight, bias=None):
    if use_native(x):
        return _nat().linear(x, weight, bias)
    return F.linear(x, weight, bias)


def cross_entropy(logits, target):
    """Mean-reduced softmax cross entropy (dpp.py:40 ``nn.CrossEntropyLoss()``)."""
    if use_native(logits):
        return _nat().cross_entropy(logits, target)
    return F.cross_entropy(logits, target)


def _nat_tf():
    from . import native_tf

    return native_tf


def layer_norm(x, weight, bias, eps=1e-5):
    """LayerNorm over the last dimension (GPT-2 blocks); hand-written kernel on CUDA (ops/native_tf.py)."""
    if use_native(x):
        return _nat_tf().layer_norm(x, weight, bias, eps)
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


def gelu(x):
    """tanh-approximated GELU, the GPT-2 activation."""
    if use_native(x):
        return _nat_tf().gelu(x)
    return F.gelu(x, approximate="tanh")


def causal_self_attention(qkv, heads: int):
    """qkv: [B, T, 3*D] fused projection output -> [B, T, D]; causal softmax(QK^T/sqrt(d))V per head.  On CUDA the
    fused tcgen05 attention kernel (csrc/kernels/attention_tcgen05.cu) reads Q, K, V in place."""
    if use_native(qkv):
        return _nat_tf().causal_self_attention(qkv, heads)
    b, t, d3 = qkv.shape
    d = d3 // 3
    q, k, v = (z.view(b, t, heads, d // heads).transpose(1, 2) for z in qkv.split(d, dim=2))
    y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return y.transpose(1, 2).reshape(b, t, d)


def embedding(idx, wte, wpe):
    """Token + position embedding: wte[idx] + wpe[0:T]."""
    if use_native(wte):
        return _nat_tf().embedding(idx, wte, wpe)
    return F.embedding(idx, wte) + wpe[: idx.shape[1]]
