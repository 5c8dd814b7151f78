"""CPU tier: the optimizers' reference paths against torch.optim, and the op fallbacks (what the same modules run when
no GPU is present) against torch.nn.functional."""
import pytest
import torch
import torch.nn.functional as F

import distributeddataparallel_b200 as d
from distributeddataparallel_b200 import ops


def _pair():
    a = [torch.nn.Parameter(torch.randn(5, 3)), torch.nn.Parameter(torch.randn(7))]
    return a, [torch.nn.Parameter(t.detach().clone()) for t in a]


@pytest.mark.parametrize("kind", ["sgd", "sgd_nesterov", "adam", "adamw"])
def test_optimizers_match_torch_on_cpu(kind):
    torch.manual_seed(0)
    a, b = _pair()
    if kind == "sgd":
        oa, ob = d.SGD(a, lr=0.1, momentum=0.9, dampening=0.1, weight_decay=1e-3), \
            torch.optim.SGD(b, lr=0.1, momentum=0.9, dampening=0.1, weight_decay=1e-3)
    elif kind == "sgd_nesterov":
        oa, ob = d.SGD(a, lr=0.1, momentum=0.9, nesterov=True, maximize=True), \
            torch.optim.SGD(b, lr=0.1, momentum=0.9, nesterov=True, maximize=True)
    elif kind == "adam":
        oa, ob = d.Adam(a, lr=1e-2, weight_decay=1e-2), torch.optim.Adam(b, lr=1e-2, weight_decay=1e-2)
    else:
        oa, ob = d.Adam(a, lr=1e-2, weight_decay=1e-2, decoupled=True), torch.optim.AdamW(b, lr=1e-2, weight_decay=1e-2)
    for _ in range(5):
        for x, y in zip(a, b):
            g = torch.randn_like(x)
            x.grad, y.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    for x, y in zip(a, b):
        assert torch.allclose(x, y, rtol=1e-6, atol=1e-7)
    # the state layout is torch's: a stock optimizer loads it
    ob.load_state_dict(oa.state_dict())
    assert len(oa.state_dict()["state"]) == len(ob.state_dict()["state"])


def test_sgd_step_skip_and_subset_api_on_cpu():
    """step(skip=ids) leaves the listed parameters alone (the hook the per-bucket optimizer path relies on)."""
    torch.manual_seed(1)
    a, _ = _pair()
    opt = d.SGD(a, lr=0.5)
    before = [p.detach().clone() for p in a]
    for p in a:
        p.grad = torch.ones_like(p)
    opt.step(skip={id(a[0])})
    assert torch.equal(a[0], before[0]) and torch.allclose(a[1], before[1] - 0.5)
    assert opt.supports_subset_steps() is False          # CPU parameters: the fused per-bucket path is CUDA-only


def test_op_fallbacks_match_torch_functional():
    torch.manual_seed(2)
    x, w, b = torch.randn(2, 8, 16), torch.randn(16), torch.randn(16)
    assert torch.equal(ops.layer_norm(x, w, b), F.layer_norm(x, (16,), w, b))
    assert torch.equal(ops.gelu(x), F.gelu(x, approximate="tanh"))
    qkv = torch.randn(2, 8, 48)
    q, k, v = (z.view(2, 8, 4, 4).transpose(1, 2) for z in qkv.split(16, dim=2))
    ref = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(2, 8, 16)
    assert torch.allclose(ops.causal_self_attention(qkv, 4), ref, atol=1e-6)
    idx, wte, wpe = torch.randint(0, 10, (2, 8)), torch.randn(10, 16), torch.randn(8, 16)
    assert torch.equal(ops.embedding(idx, wte, wpe), F.embedding(idx, wte) + wpe)
