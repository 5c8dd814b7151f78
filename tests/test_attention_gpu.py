"""Fused tcgen05 causal attention (csrc/kernels/attention_tcgen05.cu) against a plain fp32 PyTorch
reference of the same op, forward and backward, and against the six-GEMM formulation it replaces."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference(qkv, heads, dout):
    """fp32 attention on the bf16-rounded inputs; returns out and d(qkv)."""
    x = qkv.detach().float().requires_grad_(True)
    b, t, three_c = x.shape
    c = three_c // 3
    d = c // heads
    q, k, v = (z.reshape(b, t, heads, d).transpose(1, 2) for z in x.split(c, dim=2))
    s = (q @ k.transpose(-1, -2)) / math.sqrt(d)
    mask = torch.ones(t, t, device=x.device, dtype=torch.bool).tril()
    p = torch.softmax(s.masked_fill(~mask, float("-inf")), dim=-1)
    out = (p @ v).transpose(1, 2).reshape(b, t, c)
    out.backward(dout.float())
    return out.detach(), x.grad


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("b,heads,t", [(1, 1, 128), (2, 3, 256), (1, 2, 1024), (2, 2, 200), (1, 4, 77), (3, 2, 384)])
def test_fused_attention_matches_fp32_reference(b, heads, t):
    from distributeddataparallel_b200.ops import native, native_tf

    torch.manual_seed(b * 1000 + t)
    c = heads * 64
    qkv = (torch.randn(b, t, 3 * c, device="cuda") * 1.5).to(torch.bfloat16).requires_grad_(True)
    dout = torch.randn(b, t, c, device="cuda").to(torch.bfloat16)
    native_tf.set_fused_attention(True)
    out = native_tf.causal_self_attention(qkv, heads)
    out.backward(dout)
    torch.cuda.synchronize()
    assert native.tc_error_word() == 0
    ref_out, ref_grad = _reference(qkv, heads, dout)
    assert _rel(out, ref_out) < 1e-2, _rel(out, ref_out)
    c3 = qkv.grad.shape[-1] // 3
    for name, sl in (("dq", slice(0, c3)), ("dk", slice(c3, 2 * c3)), ("dv", slice(2 * c3, 3 * c3))):
        err = _rel(qkv.grad[..., sl], ref_grad[..., sl])
        assert err < 2e-2, (name, err)


def test_fused_attention_agrees_with_gemm_formulation():
    from distributeddataparallel_b200.ops import native_tf

    torch.manual_seed(3)
    b, heads, t = 2, 4, 512
    c = heads * 64
    base = torch.randn(b, t, 3 * c, device="cuda").to(torch.bfloat16)
    dout = torch.randn(b, t, c, device="cuda").to(torch.bfloat16)
    res = []
    try:
        for fused in (True, False):
            native_tf.set_fused_attention(fused)
            x = base.clone().requires_grad_(True)
            o = native_tf.causal_self_attention(x, heads)
            o.backward(dout)
            res.append((o.detach(), x.grad))
    finally:
        native_tf.set_fused_attention(True)
    assert _rel(res[0][0], res[1][0]) < 1.5e-2
    assert _rel(res[0][1], res[1][1]) < 3e-2


def test_fused_attention_repeatable_and_leaves_no_error():
    """Two calls in a row on the same input give bitwise the same forward (fixed reduction order) and
    the gradient accumulator is re-zeroed by every backward."""
    from distributeddataparallel_b200.ops import native, native_tf

    torch.manual_seed(5)
    qkv = torch.randn(2, 256, 3 * 128, device="cuda").to(torch.bfloat16)
    dout = torch.randn(2, 256, 128, device="cuda").to(torch.bfloat16)
    outs, grads = [], []
    for _ in range(2):
        x = qkv.clone().requires_grad_(True)
        o = native_tf.causal_self_attention(x, 2)
        o.backward(dout)
        outs.append(o.detach().clone())
        grads.append(x.grad.clone())
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1])
    assert _rel(grads[0], grads[1]) < 1e-3          # dQ uses floating-point atomics: order may differ
    assert native.tc_error_word() == 0
