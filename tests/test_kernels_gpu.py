"""Kernel-unit tier (1 GPU): every sm_100a kernel against a plain PyTorch fp32 reference of
the same op, on the dpp.py layer shapes including the M=32 layers and the
ragged batch of 10."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from distributeddataparallel_b200.ops import native

    native.load_library()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    native.set_tensor_cores(False)       # this module checks the exact-fp32 FFMA kernels
    yield native
    native.set_tensor_cores("auto")


def _cl(t):
    return t.contiguous(memory_format=torch.channels_last)


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


CONV_SHAPES = [
    # (N, Cin, H, Cout, R, stride, pad)
    (32, 3, 32, 64, 7, 2, 3),     # stem
    (32, 64, 8, 64, 3, 1, 1),     # layer1
    (32, 64, 8, 128, 3, 2, 1),    # layer2.0.conv1
    (32, 128, 4, 128, 3, 1, 1),
    (32, 64, 8, 128, 1, 2, 0),    # downsample
    (32, 128, 4, 256, 3, 2, 1),
    (32, 256, 2, 256, 3, 1, 1),
    (32, 256, 2, 512, 3, 2, 1),   # layer4.0.conv1 -> 1x1 maps
    (32, 512, 1, 512, 3, 1, 1),   # M = 32
    (32, 256, 2, 512, 1, 2, 0),
    (10, 64, 8, 64, 3, 1, 1),     # ragged batch
    (10, 512, 1, 512, 3, 1, 1),
    (5, 6, 9, 10, 3, 2, 1),       # odd sizes: scalar gather paths
]


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_fprop_dgrad_wgrad(nat, shape):
    n, cin, h, cout, r, stride, pad = shape
    torch.manual_seed(0)
    x = _cl(torch.randn(n, cin, h, h, device="cuda")).requires_grad_(True)
    w = _cl(torch.randn(cout, cin, r, r, device="cuda") * 0.1).requires_grad_(True)
    y = nat.conv2d(x, w, stride, pad)
    ref = F.conv2d(x.detach().double(), w.detach().double(), None, stride, pad)
    assert y.shape == ref.shape
    assert _rel(y, ref) < 1e-5
    dy = _cl(torch.randn_like(y))
    gx, gw = torch.autograd.grad(y, (x, w), dy)
    xd, wd = x.detach().double().requires_grad_(True), w.detach().double().requires_grad_(True)
    rx, rw = torch.autograd.grad(F.conv2d(xd, wd, None, stride, pad), (xd, wd), dy.double())
    assert _rel(gx, rx) < 1e-5 and _rel(gw, rw) < 1e-5
    assert gw.stride() == w.stride()


def test_conv_fused_bn_stats(nat):
    x = _cl(torch.randn(32, 64, 8, 8, device="cuda"))
    w = _cl(torch.randn(128, 64, 3, 3, device="cuda") * 0.1)
    stats = torch.zeros(256, device="cuda")
    y = nat.conv2d(x, w, 2, 1, stats)
    assert torch.allclose(stats[:128], y.sum(dim=(0, 2, 3)), rtol=1e-4, atol=1e-3)
    assert torch.allclose(stats[128:], (y * y).sum(dim=(0, 2, 3)), rtol=1e-4, atol=1e-3)


@pytest.fixture(params=["stream", "cluster"])
def bn_mode(request, nat):
    """stream: statistics from float-atomic accumulators + two-kernel backward; cluster: the one-launch deterministic
    cluster kernels (exact two-pass forward statistics, fused backward) wherever the shape qualifies."""
    old = (nat.bn_stats_mode(), nat._BN_CLUSTER_BWD)
    nat.set_bn_stats_mode("exact" if request.param == "cluster" else "fused")
    nat._BN_CLUSTER_BWD = request.param == "cluster"
    yield request.param
    nat.set_bn_stats_mode(old[0])
    nat._BN_CLUSTER_BWD = old[1]


@pytest.mark.parametrize("n,c,h", [(32, 64, 16), (32, 64, 8), (32, 128, 4), (32, 512, 1), (10, 256, 2), (4, 2048, 7)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_batchnorm_fused(nat, bn_mode, n, c, h, relu, res):
    torch.manual_seed(1)
    x = _cl(torch.randn(n, c, h, h, device="cuda") * 2 + 0.5).requires_grad_(True)
    r = _cl(torch.randn(n, c, h, h, device="cuda")).requires_grad_(True) if res else None
    g = (torch.rand(c, device="cuda") + 0.5).requires_grad_(True)
    b = torch.randn(c, device="cuda").requires_grad_(True)
    rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    nbt = torch.zeros((), dtype=torch.int64, device="cuda")
    rm2, rv2 = rm.clone(), rv.clone()
    ws = nat.BNWorkspace(c, x.device)
    out = nat.batch_norm_act(x, g, b, rm, rv, nbt, r, relu, True, 0.1, 1e-5, ws, False)
    xd = x.detach().double().requires_grad_(True)
    gd, bd = g.detach().double().requires_grad_(True), b.detach().double().requires_grad_(True)
    rd = r.detach().double().requires_grad_(True) if res else None
    ref = F.batch_norm(xd, rm2.double(), rv2.double(), gd, bd, True, 0.1, 1e-5)
    if res:
        ref = ref + rd
    if relu:
        ref = F.relu(ref)
    assert _rel(out, ref) < 1e-4
    assert int(nbt) == 1
    m = n * h * h
    var = x.detach().double().var(dim=(0, 2, 3), unbiased=m > 1)
    assert torch.allclose(rm.double(), 0.1 * x.detach().double().mean(dim=(0, 2, 3)), rtol=1e-4, atol=1e-5)
    if m > 1:
        assert torch.allclose(rv.double(), 0.9 + 0.1 * var, rtol=1e-3, atol=1e-5)
    dout = _cl(torch.randn_like(out))
    ins = (x, g, b) + ((r,) if res else ())
    rins = (xd, gd, bd) + ((rd,) if res else ())
    grads = torch.autograd.grad(out, ins, dout)
    rgrads = torch.autograd.grad(ref, rins, dout.double())
    tol = 5e-3 if m <= 64 else 5e-4
    for a, bref in zip(grads, rgrads):
        assert _rel(a, bref) < tol
    # self-cleaning accumulators: a second round gives the same answer
    out2 = nat.batch_norm_act(x, g, b, rm, rv, nbt, r, relu, True, 0.1, 1e-5, ws, False)
    assert _rel(out2, ref) < 1e-4


def test_batchnorm_cluster_kernels_are_exact_for_large_means_and_reproducible(nat):
    """|mean| >> std: E[x^2] - E[x]^2 in fp32 loses the variance; the cluster forward centres the data first.  Run
    twice: the fixed-order cluster reductions give the same bits."""
    old = (nat.bn_stats_mode(), nat._BN_CLUSTER_BWD)
    nat.set_bn_stats_mode("exact")
    nat._BN_CLUSTER_BWD = True
    try:
        torch.manual_seed(3)
        n, c, h = 32, 128, 4
        x = _cl(torch.randn(n, c, h, h, device="cuda") * 0.05 + 50.0).requires_grad_(True)
        g = (torch.rand(c, device="cuda") + 0.5).requires_grad_(True)
        b = torch.randn(c, device="cuda").requires_grad_(True)
        dout = _cl(torch.randn(n, c, h, h, device="cuda"))
        runs = []
        for _ in range(2):
            rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
            nbt = torch.zeros((), dtype=torch.int64, device="cuda")
            ws = nat.BNWorkspace(c, x.device)
            out = nat.batch_norm_act(x, g, b, rm, rv, nbt, None, True, True, 0.1, 1e-5, ws, False)
            grads = torch.autograd.grad(out, (x, g, b), dout)
            runs.append((out.detach().clone(), rm.clone(), rv.clone()) + tuple(t.clone() for t in grads))
        for a, bb in zip(runs[0], runs[1]):
            assert torch.equal(a, bb)
        xd = x.detach().double().requires_grad_(True)
        gd, bd = g.detach().double().requires_grad_(True), b.detach().double().requires_grad_(True)
        ref = F.relu(F.batch_norm(xd, None, None, gd, bd, True, 0.1, 1e-5))
        assert _rel(runs[0][0], ref) < 1e-3
        var = x.detach().double().var(dim=(0, 2, 3), unbiased=True)
        assert torch.allclose(runs[0][2].double(), 0.9 + 0.1 * var, rtol=1e-3, atol=1e-7)
        rgrads = torch.autograd.grad(ref, (xd, gd, bd), dout.double())
        for a, bref in zip(runs[0][3:], rgrads):
            assert _rel(a, bref) < 2e-3
    finally:
        nat.set_bn_stats_mode(old[0])
        nat._BN_CLUSTER_BWD = old[1]


def test_maxpool_matches_aten_including_ties(nat):
    x = _cl(F.relu(torch.randn(32, 64, 16, 16, device="cuda"))).requires_grad_(True)   # many zero ties
    y = nat.max_pool_3x3s2(x)
    xr = x.detach().clone().requires_grad_(True)
    ref = F.max_pool2d(xr, 3, 2, 1)
    assert torch.equal(y, ref)
    dy = torch.randn_like(y)
    (gx,) = torch.autograd.grad(y, x, _cl(dy))
    (rx,) = torch.autograd.grad(ref, xr, dy)
    assert torch.allclose(gx, rx, atol=1e-6)


def test_global_avgpool(nat):
    for shape in [(32, 512, 1, 1), (8, 2048, 7, 7)]:
        x = _cl(torch.randn(*shape, device="cuda")).requires_grad_(True)
        y = nat.global_avg_pool(x)
        assert torch.allclose(y, x.mean(dim=(2, 3)), atol=1e-6)
        (gx,) = torch.autograd.grad(y, x, torch.ones_like(y))
        assert torch.allclose(gx, torch.full_like(gx, 1.0 / (shape[2] * shape[3])))


@pytest.mark.parametrize("m,k,n", [(32, 512, 10), (10, 512, 10), (64, 1024, 3072), (7, 33, 5)])
def test_linear(nat, m, k, n):
    x = torch.randn(m, k, device="cuda", requires_grad=True)
    w = (torch.randn(n, k, device="cuda") * 0.05).requires_grad_(True)
    b = torch.randn(n, device="cuda", requires_grad=True)
    y = nat.linear(x, w, b)
    ref = F.linear(x.double(), w.double(), b.double())
    assert _rel(y, ref) < 1e-5
    dy = torch.randn_like(y)
    g = torch.autograd.grad(y, (x, w, b), dy)
    r = torch.autograd.grad(ref, (x, w, b), dy.double())
    for a, bb in zip(g, r):
        assert _rel(a, bb) < 1e-5


@pytest.mark.parametrize("b,v", [(32, 10), (10, 10), (64, 1000), (8, 50257)])
def test_cross_entropy(nat, b, v):
    z = (torch.randn(b, v, device="cuda") * 3).requires_grad_(True)
    t = torch.randint(0, v, (b,), device="cuda")
    loss = nat.cross_entropy(z, t)
    zr = z.detach().double().requires_grad_(True)
    ref = F.cross_entropy(zr, t)
    assert abs(float(loss) - float(ref)) < 1e-4 * max(1.0, abs(float(ref)))
    (g,) = torch.autograd.grad(loss * 2.0, z)
    (r,) = torch.autograd.grad(ref * 2.0, zr)
    assert _rel(g, r) < 1e-4


def test_preprocess_matches_torchvision_arithmetic(nat):
    u = torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, device="cuda")
    out = nat.preprocess_u8(u)
    # torchvision runs ToTensor/Normalize on the host with true divisions; the CUDA scalar-div
    # shortcut (multiply by reciprocal) is NOT the reference arithmetic
    uc = u.cpu()
    ref = (uc.float().div(255) - torch.tensor(0.5)).div(torch.tensor(0.5))
    assert out.shape == (32, 3, 32, 32) and out.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(out.permute(0, 2, 3, 1).cpu(), ref)


@pytest.mark.parametrize("momentum,wd,nesterov", [(0.0, 0.0, False), (0.9, 1e-4, False), (0.9, 0.0, True)])
def test_fused_sgd_matches_torch(nat, momentum, wd, nesterov):
    import distributeddataparallel_b200 as d

    torch.manual_seed(0)
    shapes = [(64, 3, 7, 7), (64,), (128, 64, 3, 3), (10, 512), (5000,), (3,)]
    ps = [torch.randn(s, device="cuda").requires_grad_(True) for s in shapes]
    ps[2] = _cl(ps[2].detach()).requires_grad_(True)
    qs = [p.detach().clone().requires_grad_(True) for p in ps]
    a = d.SGD(ps, lr=0.05, momentum=momentum, weight_decay=wd, nesterov=nesterov)
    b = torch.optim.SGD(qs, lr=0.05, momentum=momentum, weight_decay=wd, nesterov=nesterov)
    for _ in range(3):
        for p, q in zip(ps, qs):
            g = torch.randn_like(p)
            p.grad, q.grad = g.clone(memory_format=torch.preserve_format), g.clone()
        a.step()
        b.step()
    for p, q in zip(ps, qs):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)
    assert a.state_dict()["param_groups"][0].keys() == b.state_dict()["param_groups"][0].keys()


@pytest.mark.parametrize("momentum", [0.0, 0.9])
def test_fused_sgd_bf16_parameters(nat, momentum):
    """bf16 parameters / gradients (vectorised path for 8-element-aligned tensors, scalar path otherwise) against
    the same update computed in fp32 on the bf16-rounded values; momentum state is fp32."""
    import distributeddataparallel_b200 as d

    torch.manual_seed(1)
    shapes = [(1024, 1024), (4096,), (50257, 16), (77,), (3, 5)]
    ps = [torch.randn(s, device="cuda").to(torch.bfloat16).requires_grad_(True) for s in shapes]
    ref = [p.detach().float() for p in ps]
    bufs = [torch.zeros_like(r) for r in ref]
    opt = d.SGD(ps, lr=0.05, momentum=momentum)
    for it in range(3):
        for p, r, b in zip(ps, ref, bufs):
            g = torch.randn_like(p)
            p.grad = g
            gf = g.float()
            if momentum:
                b.copy_(gf if it == 0 else momentum * b + gf)
                gf = b
            r.copy_((r - 0.05 * gf).to(torch.bfloat16).float())
        opt.step()
    for p, r in zip(ps, ref):
        assert torch.allclose(p.float(), r, rtol=1e-2, atol=1e-2), (p.shape, (p.float() - r).abs().max())


def test_fused_adam_matches_torch(nat):
    import distributeddataparallel_b200 as d

    torch.manual_seed(0)
    ps = [torch.randn(s, device="cuda").requires_grad_(True) for s in [(300, 7), (64,), (4097,)]]
    qs = [p.detach().clone().requires_grad_(True) for p in ps]
    a = d.Adam(ps, lr=1e-2, weight_decay=0.01)
    b = torch.optim.Adam(qs, lr=1e-2, weight_decay=0.01)
    for _ in range(4):
        for p, q in zip(ps, qs):
            g = torch.randn_like(p)
            p.grad, q.grad = g.clone(), g.clone()
        a.step()
        b.step()
    for p, q in zip(ps, qs):
        assert torch.allclose(p, q, rtol=1e-4, atol=1e-5)
    assert float(a.state[ps[0]]["step"]) == 4.0
