"""GPT-2 module: CPU plumbing (always) and native-kernel numerics vs float64 (GPU)."""
import pytest
import torch

import distributeddataparallel_b200 as d
from distributeddataparallel_b200.models.gpt2 import build_gpt2


def test_gpt2_medium_parameter_count():
    import torch.nn as nn

    with torch.device("meta"):
        m = build_gpt2("gpt2-medium")
    n = sum(p.numel() for p in m.parameters())
    assert abs(n - 354.8e6) / 354.8e6 < 0.01, n


def test_gpt2_tiny_trains_on_cpu():
    torch.manual_seed(0)
    m = build_gpt2("gpt2-tiny", seq_len=32)
    opt = d.Adam(m.parameters(), lr=3e-3)
    idx = torch.randint(0, 512, (4, 32))
    tgt = torch.roll(idx, -1, dims=1)
    first = None
    for _ in range(20):
        opt.zero_grad()
        loss = m(idx, tgt)
        loss.backward()
        opt.step()
        first = first if first is not None else float(loss)
    assert float(loss) < first - 0.5


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-12))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 5e-3), (torch.bfloat16, 6e-2)])
def test_gpt2_native_matches_fp64(dtype, tol):
    torch.manual_seed(0)
    ref = build_gpt2("gpt2-tiny", seq_len=64).double()
    ref.wte.weight.data.mul_(5.0)                  # non-trivial logits
    nat = build_gpt2("gpt2-tiny", seq_len=64)
    nat.load_state_dict(ref.state_dict())
    nat = nat.to("cuda", dtype)
    idx = torch.randint(0, 512, (4, 64))
    tgt = torch.randint(0, 512, (4, 64))
    d.ops.set_native(False)
    lr = ref(idx, tgt)
    lr.backward()
    d.ops.set_native(True)
    ln = nat(idx.cuda(), tgt.cuda())
    ln.backward()
    torch.cuda.synchronize()
    assert abs(float(ln) - float(lr)) < tol * max(1.0, abs(float(lr))), (float(ln), float(lr))
    errs = {n: _rel(p.grad.cpu(), q.grad) for (n, p), (_, q) in zip(nat.named_parameters(), ref.named_parameters())}
    worst = max(errs, key=errs.get)
    assert errs[worst] < 8 * tol, (worst, errs[worst])
    from distributeddataparallel_b200.ops import native

    assert native.tc_error_word() == 0


@pytest.mark.gpu
def test_gpt2_ddp_graph_step_bf16():
    torch.manual_seed(0)
    model = build_gpt2("gpt2-tiny", seq_len=64).to("cuda", torch.bfloat16)
    ddp = d.DDP(model, device_ids=[0])
    opt = d.SGD(ddp.parameters(), lr=0.05)
    step = d.TrainStep(ddp, None, opt, cuda_graph=True)
    idx = torch.randint(0, 512, (4, 64), device="cuda")
    tgt = torch.roll(idx, -1, dims=1)
    losses = [float(step(idx, tgt)) for _ in range(12)]
    assert step.captured and step.replays >= 8
    assert losses[-1] < losses[0] - 0.2, losses


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 5e-3), (torch.bfloat16, 4e-2)])
def test_causal_attention_multi_tile_matches_sdpa(dtype, tol):
    """T = 384 spans three 128-row tiles: exercises the causal tile skipping of all six attention GEMMs."""
    from distributeddataparallel_b200.ops import native_tf

    torch.manual_seed(0)
    b, t, h, dh = 2, 384, 4, 64
    qkv64 = torch.randn(b, t, 3 * h * dh, dtype=torch.float64)
    qkv = qkv64.to("cuda", dtype).requires_grad_(True)
    out = native_tf.causal_self_attention(qkv, h)
    ref_in = qkv64.clone().requires_grad_(True)
    q, k, v = ref_in.view(b, t, 3, h, dh).permute(2, 0, 3, 1, 4)
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(b, t, h * dh)
    assert _rel(out.cpu(), ref) < tol, _rel(out.cpu(), ref)
    g = torch.randn(b, t, h * dh, dtype=torch.float64)
    out.backward(g.to("cuda", dtype))
    ref.backward(g)
    assert _rel(qkv.grad.cpu(), ref_in.grad) < 2 * tol, _rel(qkv.grad.cpu(), ref_in.grad)


@pytest.mark.gpu
@pytest.mark.parametrize("m,dm", [(2048, 256), (4096, 1024)])
def test_fused_mlp_epilogues_match_fp32_and_unfused(m, dm):
    """GELU / GELU' / residual-add epilogues of the persistent GEMM vs the separate kernels and an fp32 reference."""
    from distributeddataparallel_b200.ops import native, native_tf

    torch.manual_seed(11)
    f = 4 * dm
    mk = lambda *s, sc=1.0: (torch.randn(*s, device="cuda") * sc).to(torch.bfloat16)
    x, res, dy = mk(m, dm), mk(m, dm), mk(m, dm)
    w1, b1, w2, b2 = mk(f, dm, sc=dm ** -0.5), mk(f, sc=0.1), mk(dm, f, sc=f ** -0.5), mk(dm, sc=0.1)
    outs = {}
    try:
        for fused in (True, False):
            native_tf.set_fused_epilogues(fused)
            leaves = [t.clone().requires_grad_(True) for t in (x, w1, b1, w2, b2, res)]
            y = native_tf.mlp(*leaves)
            y.backward(dy)
            outs[fused] = [y.detach()] + [t.grad for t in leaves]
    finally:
        native_tf.set_fused_epilogues(True)
    torch.cuda.synchronize()
    assert native.tc_error_word() == 0
    leaves = [t.float().requires_grad_(True) for t in (x, w1, b1, w2, b2, res)]
    xr, w1r, b1r, w2r, b2r, rr = leaves
    yr = rr + torch.nn.functional.gelu(xr @ w1r.t() + b1r, approximate="tanh") @ w2r.t() + b2r
    yr.backward(dy.float())
    ref = [yr.detach()] + [t.grad for t in leaves]
    names = ["y", "dx", "dw1", "db1", "dw2", "db2", "dres"]
    for name, a, b_, r in zip(names, outs[True], outs[False], ref):
        assert _rel(a, r) < 2e-2, (name, "fused vs fp32", _rel(a, r))
        assert _rel(b_, r) < 2e-2, (name, "unfused vs fp32", _rel(b_, r))


@pytest.mark.gpu
def test_linear_residual_epilogue():
    from distributeddataparallel_b200.ops import native_tf

    torch.manual_seed(12)
    m, k, n = 4096, 512, 1024
    x = torch.randn(m, k, device="cuda").to(torch.bfloat16)
    w = (torch.randn(n, k, device="cuda") * k ** -0.5).to(torch.bfloat16)
    b = torch.randn(n, device="cuda").to(torch.bfloat16)
    r = torch.randn(m, n, device="cuda").to(torch.bfloat16).requires_grad_(True)
    y = native_tf.linear_tc(x, w, b, residual=r)
    y.backward(torch.ones_like(y))
    ref = x.float() @ w.float().t() + b.float() + r.detach().float()
    assert _rel(y, ref) < 1e-2
    assert torch.equal(r.grad, torch.ones_like(r))


@pytest.mark.gpu
@pytest.mark.parametrize("d_model", [768, 1024, 1280, 512])
def test_layernorm_with_residual_branch_and_one_launch_bias_grad(d_model):
    """LayerNorm whose backward also absorbs the residual branch's gradient (fused widths 768/1024/1280, generic
    kernels otherwise), and the one-launch bf16 bias gradient, against fp32 references; run twice to check the
    zero-on-exit accumulators."""
    from distributeddataparallel_b200.ops import native_tf

    torch.manual_seed(21)
    m = 2048
    for _ in range(2):
        x = torch.randn(m, d_model, device="cuda").to(torch.bfloat16).requires_grad_(True)
        g = (1 + 0.1 * torch.randn(d_model, device="cuda")).to(torch.bfloat16).requires_grad_(True)
        b = (0.1 * torch.randn(d_model, device="cuda")).to(torch.bfloat16).requires_grad_(True)
        w = (torch.randn(d_model, d_model, device="cuda") * d_model ** -0.5).to(torch.bfloat16).requires_grad_(True)
        bias = torch.zeros(d_model, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        dy = torch.randn(m, d_model, device="cuda").to(torch.bfloat16)
        h, xr = native_tf.layer_norm_res(x, g, b, 1e-5)
        out = native_tf.linear_tc(h, w, bias, residual=xr)
        out.backward(dy)
        xf, gf, bf_, wf, biasf = (t.detach().float().requires_grad_(True) for t in (x, g, b, w, bias))
        ref = torch.nn.functional.layer_norm(xf, (d_model,), gf, bf_, 1e-5) @ wf.t() + biasf + xf
        ref.backward(dy.float())
        assert _rel(out, ref) < 1e-2
        for name, a, r in (("dx", x.grad, xf.grad), ("dgamma", g.grad, gf.grad), ("dbeta", b.grad, bf_.grad),
                           ("dbias", bias.grad, biasf.grad)):
            assert _rel(a, r) < 2e-2, (name, _rel(a, r))
