"""Autograd Functions for the transformer path (GPT-2 configuration): every matrix product is
the tcgen05 GEMM (``b200_gemm_tc_ex``; TF32 math on fp32 tensors, bf16 x bf16 -> fp32 accumulate
on bf16 tensors), LayerNorm / GELU / causal softmax / embedding / cross-entropy are the kernels
in ``csrc/kernels/transformer.cu``.  Attention is computed per (batch, head) with the *batched*
GEMM reading Q, K, V straight out of the fused QKV projection buffer (no permute copies).

SURVEY.md W17 / BASELINE.json config 4 (the reference itself has no transformer: dpp.py trains
a CNN).
"""
from __future__ import annotations

import ctypes
import math
import os
from typing import Optional

import torch

from . import native
from .native import _call, _i, _ptr, _stream, _sz, _grad_target, scratch, _ws_args

_ll11 = ctypes.c_longlong * 11


def _bf(t: torch.Tensor) -> int:
    return int(t.dtype == torch.bfloat16)


def gemm_ex(a, b, out, mode: int, m: int, n: int, k: int, lda: int, ldb: int, ldo: int,
            bias: Optional[torch.Tensor] = None, batch: int = 1, inner: int = 1,
            strides=(0, 0, 0, 0, 0, 0), accumulate: bool = False, causal: int = 0) -> None:
    """out = op(a) @ op(b) on the tcgen05 kernel (see b200_gemm_tc_ex for modes / descriptors)."""
    assert a.dtype == b.dtype and a.dtype in (torch.float32, torch.bfloat16), (a.dtype, b.dtype)
    flags = (_bf(a) | (_bf(out) << 1) | ((_bf(bias) if bias is not None else 0) << 2) | (int(accumulate) << 3)
             | (int(causal) << 4))
    desc = _ll11(lda, ldb, ldo, batch, inner, *strides)
    s = scratch(a.device)
    _call("b200_gemm_tc_ex", _ptr(a), _ptr(b), _ptr(out), _ptr(bias), _i(m), _i(n), _i(k), _i(mode), _i(flags),
          desc, *_ws_args(a.device), _ptr(s.error), _stream())


# ---------------------------------------------------------------------------------------------
_FUSED_EPILOGUES = os.environ.get("B200DDP_FUSED_EPILOGUES", "1") != "0"


def set_fused_epilogues(enabled: bool) -> None:
    global _FUSED_EPILOGUES
    _FUSED_EPILOGUES = bool(enabled)


def gemm_epi(a, b, out, mode: int, m: int, n: int, k: int, lda: int, ldb: int, ldo: int, epi: int,
             bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None,
             aux_out: Optional[torch.Tensor] = None) -> bool:
    """GEMM with a fused epilogue on the persistent 2-CTA kernel (``b200_gemm_tc_epi``): epi 1 GELU forward
    (``aux_out`` <- pre-activation), 2 GELU backward (times gelu'(``aux``)), 3 residual add (+ ``aux``).
    Returns False when the shape is not eligible (nothing was launched)."""
    if not _FUSED_EPILOGUES or a.dtype != torch.bfloat16 or out.dtype != torch.bfloat16:
        return False
    flags = 1 | 2 | ((_bf(bias) if bias is not None else 0) << 2)
    desc = _ll11(lda, ldb, ldo, 1, 1, 0, 0, 0, 0, 0, 0)
    s = scratch(a.device)
    return _try_call("b200_gemm_tc_epi", _ptr(a), _ptr(b), _ptr(out), _ptr(bias), _ptr(aux), _ptr(aux_out), _i(epi),
                     _i(m), _i(n), _i(k), _i(mode), _i(flags), desc, *_ws_args(a.device), _ptr(s.error), _stream()) == 0


def _bias_grad(dy2, param, slot):
    m, n = dy2.shape
    db, _ = _grad_target(param, slot, param)
    sc = scratch(dy2.device)
    if _bf(dy2) and db.dtype == torch.bfloat16 and m >= 1024 and n % 8 == 0 and n <= sc.red_acc.numel() and db.is_contiguous():
        _call("b200_colsum_bf16_fused", _ptr(dy2), _ptr(db), _ptr(sc.red_acc), _ptr(sc.red_ticket), _i(m), _i(n), _stream())
        return db
    if _bf(dy2) and m >= 1024 and n % 2 == 0:
        acc = torch.zeros(n, dtype=torch.float32, device=dy2.device)
        _call("b200_colsum_bf16_acc", _ptr(dy2), _ptr(acc), _i(m), _i(n), _stream())
        _call("b200_cast_from_f32", _ptr(acc), _ptr(db), _sz(n), _i(0), _i(1), _stream())
    elif _bf(dy2):
        _call("b200_colsum_bf16", _ptr(dy2), _ptr(db), _i(m), _i(n), _i(0), _stream())
    else:
        _call("b200_colsum", _ptr(dy2), _ptr(db), _i(m), _i(n), _i(0), _stream())
    return db


def _weight_grad(dy2, x2, w, param, slot):
    (m, k), n = x2.shape, w.shape[0]
    dw, _ = _grad_target(param, slot, w)
    gemm_ex(dy2, x2, dw, 2, n, k, m, n, k, k)                           # dW = dy^T @ x
    return dw


class _LinearTC(torch.autograd.Function):
    """y = x W^T + b (+ residual): the residual add rides in the GEMM epilogue when the shape allows."""

    @staticmethod
    def forward(ctx, x, w, b, residual=None):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        m, k = x2.shape
        n = w.shape[0]
        y = torch.empty((m, n), dtype=x.dtype, device=x.device)
        fused = False
        if residual is not None:
            r2 = residual.reshape(m, n)
            fused = r2.is_contiguous() and gemm_epi(x2, w, y, 0, m, n, k, k, k, n, 3, bias=b, aux=r2)
        if not fused:
            gemm_ex(x2, w, y, 0, m, n, k, k, k, n, bias=b)
            if residual is not None:
                y = y + residual.reshape(m, n)
        ctx.save_for_backward(x2, w)
        ctx.lead = x.shape[:-1]
        ctx.params = (w, b)
        ctx.slots = (getattr(w, "_grad_slot", None), getattr(b, "_grad_slot", None) if b is not None else None)
        return y.reshape(*ctx.lead, n)

    @staticmethod
    def backward(ctx, dy):
        x2, w = ctx.saved_tensors
        m, k = x2.shape
        n = w.shape[0]
        dy2 = dy.reshape(m, n)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = dres = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((m, k), dtype=x2.dtype, device=dy.device)
            gemm_ex(dy2, w, dx, 1, m, k, n, n, k, k)                    # dx = dy @ W
            dx = dx.reshape(*ctx.lead, k)
        if ctx.needs_input_grad[1]:
            dw = _weight_grad(dy2, x2, w, ctx.params[0], ctx.slots[0])
        if ctx.params[1] is not None and ctx.needs_input_grad[2]:
            db = _bias_grad(dy2, ctx.params[1], ctx.slots[1])
        if len(ctx.needs_input_grad) > 3 and ctx.needs_input_grad[3]:
            dres = dy
        return dx, dw, db, dres


def linear_tc(x, w, b=None, residual=None):
    return _LinearTC.apply(x, w, b, residual)


class _MLP(torch.autograd.Function):
    """y = residual + gelu(x W1^T + b1) W2^T + b2 with GELU, its derivative and the residual add fused into
    the epilogues of the three GEMMs they follow (transformer block feed-forward)."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        m, d = x2.shape
        f = w1.shape[0]
        pre = torch.empty((m, f), dtype=x.dtype, device=x.device)
        act = torch.empty_like(pre)
        if not gemm_epi(x2, w1, act, 0, m, f, d, d, d, f, 1, bias=b1, aux_out=pre):
            gemm_ex(x2, w1, pre, 0, m, f, d, d, d, f, bias=b1)
            _call("b200_gelu_fwd", _ptr(pre), _ptr(act), _sz(pre.numel()), _i(_bf(pre)), _stream())
        y = torch.empty((m, d), dtype=x.dtype, device=x.device)
        r2 = residual.reshape(m, d)
        if not (r2.is_contiguous() and gemm_epi(act, w2, y, 0, m, d, f, f, f, d, 3, bias=b2, aux=r2)):
            gemm_ex(act, w2, y, 0, m, d, f, f, f, d, bias=b2)
            y = y + r2
        ctx.save_for_backward(x2, pre, act, w1, w2)
        ctx.shape = x.shape
        ctx.params = (w1, b1, w2, b2)
        ctx.slots = tuple(getattr(p, "_grad_slot", None) for p in ctx.params)
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, pre, act, w1, w2 = ctx.saved_tensors
        m, d = x2.shape
        f = w1.shape[0]
        dy2 = dy.reshape(m, d)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dpre = torch.empty_like(pre)
        if not gemm_epi(dy2, w2, dpre, 1, m, f, d, d, f, f, 2, aux=pre):       # (dy W2) * gelu'(pre)
            gemm_ex(dy2, w2, dpre, 1, m, f, d, d, f, f)
            _call("b200_gelu_bwd", _ptr(pre), _ptr(dpre), _ptr(dpre), _sz(pre.numel()), _i(_bf(pre)), _stream())
        dw2 = _weight_grad(dy2, act, w2, ctx.params[2], ctx.slots[2])
        db2 = _bias_grad(dy2, ctx.params[3], ctx.slots[3])
        dx = torch.empty((m, d), dtype=x2.dtype, device=dy.device)
        gemm_ex(dpre, w1, dx, 1, m, d, f, f, d, d)
        dw1 = _weight_grad(dpre, x2, w1, ctx.params[0], ctx.slots[0])
        db1 = _bias_grad(dpre, ctx.params[1], ctx.slots[1])
        return dx.reshape(ctx.shape), dw1, db1, dw2, db2, dy


def mlp(x, w1, b1, w2, b2, residual):
    return _MLP.apply(x, w1, b1, w2, b2, residual)


# ---------------------------------------------------------------------------------------------
class _LayerNorm(torch.autograd.Function):
    """y = LayerNorm(x).  With ``with_residual`` the Function also hands back an alias of ``x`` for the residual
    branch, so that branch's gradient comes back into this backward and is added inside the kernel."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, with_residual=False):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        m, d = x2.shape
        y = torch.empty_like(x2)
        mean = torch.empty(m, dtype=torch.float32, device=x.device)
        rstd = torch.empty(m, dtype=torch.float32, device=x.device)
        _call("b200_layernorm_fwd", _ptr(x2), _ptr(gamma), _ptr(beta), _ptr(y), _ptr(mean), _ptr(rstd), _i(m), _i(d),
              ctypes.c_float(eps), _i(_bf(x2)), _stream())
        ctx.save_for_backward(x2, gamma, mean, rstd)
        ctx.shape = x.shape
        ctx.params = (gamma, beta)
        ctx.slots = (getattr(gamma, "_grad_slot", None), getattr(beta, "_grad_slot", None))
        ctx.with_residual = with_residual
        if with_residual:
            return y.reshape(x.shape), x2.view(x.shape)
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy, dres=None):
        x2, gamma, mean, rstd = ctx.saved_tensors
        m, d = x2.shape
        dy2 = dy.reshape(m, d).contiguous()
        dx = torch.empty_like(x2)
        dgamma, _ = _grad_target(ctx.params[0], ctx.slots[0], gamma)
        dbeta, _ = _grad_target(ctx.params[1], ctx.slots[1], gamma)
        if dres is not None:
            dres = dres.reshape(m, d)
            if not dres.is_contiguous():
                dres = dres.contiguous()
        sc = scratch(dy.device)
        fused = (_bf(x2) and gamma.dtype == torch.bfloat16 and dgamma.is_contiguous() and dbeta.is_contiguous()
                 and _try_call("b200_layernorm_bwd_fused", _ptr(dy2), _ptr(x2), _ptr(gamma), _ptr(mean), _ptr(rstd), _ptr(dres),
                               _ptr(dx), _ptr(dgamma), _ptr(dbeta), _ptr(sc.red_acc), _ptr(sc.red_ticket[1:]), _i(m), _i(d),
                               _stream()) == 0)
        if not fused:
            acc = torch.zeros(2 * d, dtype=torch.float32, device=dy.device)
            _call("b200_layernorm_bwd", _ptr(dy2), _ptr(x2), _ptr(gamma), _ptr(mean), _ptr(rstd), _ptr(dx), _ptr(acc),
                  _ptr(acc[d:]), _i(m), _i(d), _i(_bf(x2)), _stream())
            _call("b200_cast_from_f32", _ptr(acc), _ptr(dgamma), _sz(d), _i(0), _i(_bf(gamma)), _stream())
            _call("b200_cast_from_f32", _ptr(acc[d:]), _ptr(dbeta), _sz(d), _i(0), _i(_bf(gamma)), _stream())
            if dres is not None:
                dx = dx + dres
        return dx.reshape(ctx.shape), dgamma, dbeta, None, None


def layer_norm(x, weight, bias, eps=1e-5):
    return _LayerNorm.apply(x, weight, bias, float(eps))


def layer_norm_res(x, weight, bias, eps=1e-5):
    """(LayerNorm(x), x): use the second result for the residual connection around the normalised branch --
    its gradient is then added to the LayerNorm input gradient inside the backward kernel."""
    return _LayerNorm.apply(x, weight, bias, float(eps), True)


class _Gelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xc = x.contiguous()
        y = torch.empty_like(xc)
        _call("b200_gelu_fwd", _ptr(xc), _ptr(y), _sz(xc.numel()), _i(_bf(xc)), _stream())
        ctx.save_for_backward(xc)
        return y

    @staticmethod
    def backward(ctx, dy):
        (xc,) = ctx.saved_tensors
        dyc = dy.contiguous()
        dx = torch.empty_like(xc)
        _call("b200_gelu_bwd", _ptr(xc), _ptr(dyc), _ptr(dx), _sz(xc.numel()), _i(_bf(xc)), _stream())
        return dx


def gelu(x):
    return _Gelu.apply(x)


# ---------------------------------------------------------------------------------------------
_FUSED_ATTENTION = os.environ.get("B200DDP_FUSED_ATTENTION", "1") != "0"


def set_fused_attention(enabled: bool) -> None:
    global _FUSED_ATTENTION
    _FUSED_ATTENTION = bool(enabled)


def _try_call(name: str, *args) -> int:
    """Like ``_call`` but hands 801 (shape / descriptor not supported) back to the caller."""
    rc = getattr(native.load_library(), name)(*args)
    if rc not in (0, 801):
        raise RuntimeError(f"{name} failed: CUDA error {rc}")
    if rc == 0:
        native._launches += 1
    return rc


class _CausalSelfAttention(torch.autograd.Function):
    """qkv: [B, T, 3*H*D] (the fused projection output)  ->  [B, T, H*D]."""

    @staticmethod
    def forward(ctx, qkv, heads):
        qkv = qkv.contiguous()
        b, t, three_hd = qkv.shape
        hd = three_hd // 3
        d = hd // heads
        z = b * heads
        scale = 1.0 / math.sqrt(d)
        ctx.fused = False
        if _FUSED_ATTENTION and d == 64 and qkv.dtype == torch.bfloat16:
            # one kernel: scores stay in TMEM / shared memory (csrc/kernels/attention_tcgen05.cu)
            out = torch.empty((b, t, hd), dtype=qkv.dtype, device=qkv.device)
            lse = torch.empty((z, t), dtype=torch.float32, device=qkv.device)
            rc = _try_call("b200_attention_fwd", _ptr(qkv), _ptr(out), _ptr(lse), _i(b), _i(heads), _i(t),
                           _ptr(scratch(qkv.device).error), _stream())
            if rc == 0:
                ctx.fused = True
                ctx.save_for_backward(qkv, out, lse)
                ctx.cfg = (b, t, heads, d, scale)
                return out
        es = qkv.element_size()
        q, k, v = qkv.view(-1)[0:], qkv.view(-1)[hd:], qkv.view(-1)[2 * hd:]      # same buffer, column offsets
        p = torch.empty((z, t, t), dtype=qkv.dtype, device=qkv.device)
        row = 3 * hd                                                              # pitch of one token's qkv row
        # S[z] = Q[z] K[z]^T      (batch z = (b, h): offset b*T*row + h*D)
        gemm_ex(q, k, p, 0, t, t, d, row, row, t, batch=z, inner=heads,
                strides=(t * row, d, t * row, d, heads * t * t, t * t), causal=1)
        _call("b200_causal_softmax_fwd", _ptr(p), _i(t), _sz(z * t), ctypes.c_float(scale), _i(_bf(p)), _stream())
        out = torch.empty((b, t, hd), dtype=qkv.dtype, device=qkv.device)
        # O[z] = P[z] V[z]        (written into [B, T, H, D] layout: pitch hd, offset b*T*hd + h*D)
        gemm_ex(p, v, out, 1, t, d, t, t, row, hd, batch=z, inner=heads,
                strides=(heads * t * t, t * t, t * row, d, t * hd, d), causal=2)
        ctx.save_for_backward(qkv, p)
        ctx.cfg = (b, t, heads, d, scale)
        return out

    @staticmethod
    def backward(ctx, dout):
        b, t, heads, d, scale = ctx.cfg
        if ctx.fused:
            qkv, out, lse = ctx.saved_tensors
            dout = dout.contiguous()
            dqkv = torch.empty_like(qkv)
            delta = torch.empty_like(lse)
            dq32 = torch.empty((b * t, heads * d), dtype=torch.float32, device=qkv.device)
            _call("b200_attention_bwd", _ptr(qkv), _ptr(out), _ptr(dout), _ptr(lse), _ptr(delta), _ptr(dq32), _ptr(dqkv),
                  _i(b), _i(heads), _i(t), _ptr(scratch(qkv.device).error), _stream(), kernels=4)
            return dqkv, None
        qkv, p = ctx.saved_tensors
        hd, z, row = heads * d, b * heads, 3 * heads * d
        dout = dout.contiguous()
        q, k, v = qkv.view(-1)[0:], qkv.view(-1)[hd:], qkv.view(-1)[2 * hd:]
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv.view(-1)[0:], dqkv.view(-1)[hd:], dqkv.view(-1)[2 * hd:]
        sq = (heads * t * t, t * t)
        # dV[z] = P^T dO
        gemm_ex(p, dout, dv, 2, t, d, t, t, hd, row, batch=z, inner=heads, strides=(*sq, t * hd, d, t * row, d),
                causal=3)
        # dP[z] = dO V^T
        dp = torch.empty_like(p)
        gemm_ex(dout, v, dp, 0, t, t, d, hd, row, t, batch=z, inner=heads, strides=(t * hd, d, t * row, d, *sq),
                causal=1)
        _call("b200_causal_softmax_bwd", _ptr(p), _ptr(dp), _i(t), _sz(z * t), ctypes.c_float(scale), _i(_bf(p)),
              _stream())
        # dQ[z] = dS K ;  dK[z] = dS^T Q
        gemm_ex(dp, k, dq, 1, t, d, t, t, row, row, batch=z, inner=heads, strides=(*sq, t * row, d, t * row, d),
                causal=2)
        gemm_ex(dp, q, dk, 2, t, d, t, t, row, row, batch=z, inner=heads, strides=(*sq, t * row, d, t * row, d),
                causal=3)
        return dqkv, None


def causal_self_attention(qkv, heads: int):
    return _CausalSelfAttention.apply(qkv, int(heads))


# ---------------------------------------------------------------------------------------------
class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, idx, wte, wpe):
        b, t = idx.shape
        d = wte.shape[1]
        out = torch.empty((b, t, d), dtype=wte.dtype, device=wte.device)
        idx = idx.contiguous()
        _call("b200_embedding_fwd", _ptr(idx), _ptr(wte), _ptr(wpe), _ptr(out), _sz(b * t), _i(t), _i(d),
              _i(_bf(wte)), _stream())
        ctx.save_for_backward(idx)
        ctx.params = (wte, wpe)
        ctx.slots = (getattr(wte, "_grad_slot", None), getattr(wpe, "_grad_slot", None))
        return out

    @staticmethod
    def backward(ctx, dout):
        (idx,) = ctx.saved_tensors
        wte, wpe = ctx.params
        b, t = idx.shape
        d = wte.shape[1]
        dout = dout.contiguous()
        acc_te = torch.zeros(wte.shape, dtype=torch.float32, device=dout.device)
        acc_pe = torch.zeros(wpe.shape, dtype=torch.float32, device=dout.device)
        _call("b200_embedding_bwd", _ptr(idx), _ptr(dout), _ptr(acc_te), _ptr(acc_pe), _sz(b * t), _i(t), _i(d),
              _i(_bf(dout)), _stream())
        if wte.dtype == torch.float32:
            dwte, dwpe = acc_te, acc_pe
        else:
            dwte, dwpe = torch.empty_like(wte), torch.empty_like(wpe)
            _call("b200_cast_from_f32", _ptr(acc_te), _ptr(dwte), _sz(acc_te.numel()), _i(0), _i(1), _stream())
            _call("b200_cast_from_f32", _ptr(acc_pe), _ptr(dwpe), _sz(acc_pe.numel()), _i(0), _i(1), _stream())
        return None, dwte, dwpe


def embedding(idx, wte, wpe):
    return _Embedding.apply(idx, wte, wpe)


# ---------------------------------------------------------------------------------------------
class _LMHeadLoss(torch.autograd.Function):
    """loss = mean CE(x @ wte^T, target): logits GEMM + in-place softmax/CE + both backward GEMMs."""

    @staticmethod
    def forward(ctx, x, wte, target):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        m, d = x2.shape
        v = wte.shape[0]
        vec = 8 if x2.dtype == torch.bfloat16 else 4
        ld = -(-v // vec) * vec                              # padded pitch keeps rows 16-byte aligned
        logits = torch.empty((m, ld), dtype=x2.dtype, device=x.device)
        gemm_ex(x2, wte, logits, 0, m, v, d, d, d, ld)
        loss = torch.empty((), dtype=torch.float32, device=x.device)
        row_loss = torch.empty(m, dtype=torch.float32, device=x.device)
        tgt = target.reshape(-1).contiguous()
        # dlogits overwrite the logits (the kernel reads each element before writing it)
        _call("b200_xent_rows", _ptr(logits), _ptr(tgt), _ptr(logits), _ptr(row_loss), _ptr(loss), _i(m), _i(v),
              _i(ld), _i(_bf(logits)), _stream(), kernels=2)
        ctx.save_for_backward(x2, wte, logits)
        ctx.cfg = (m, d, v, ld, x.shape)
        ctx.param = wte
        ctx.slot = getattr(wte, "_grad_slot", None)
        return loss

    @staticmethod
    def backward(ctx, gout):
        x2, wte, dlogits = ctx.saved_tensors
        m, d, v, ld, xshape = ctx.cfg
        dx = torch.empty((m, d), dtype=x2.dtype, device=x2.device)
        gemm_ex(dlogits, wte, dx, 1, m, d, v, ld, d, d)           # dx = dlogits @ wte
        dw = torch.empty_like(wte)                                # (tied with the embedding: autograd sums)
        gemm_ex(dlogits, x2, dw, 2, v, d, m, ld, d, d)            # dwte = dlogits^T @ x
        # the incoming loss gradient scales the two (small) results instead of the [M, V] dlogits
        g = gout.contiguous().float()
        _call("b200_scale_inplace", _ptr(dx), _sz(dx.numel()), _ptr(g), _i(_bf(dx)), _stream())
        _call("b200_scale_inplace", _ptr(dw), _sz(dw.numel()), _ptr(g), _i(_bf(dw)), _stream())
        return dx.reshape(xshape), dw, None


def lm_head_loss(x, wte, target):
    return _LMHeadLoss.apply(x, wte, target)
