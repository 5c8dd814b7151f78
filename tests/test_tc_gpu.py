"""tcgen05 / TMEM kernels vs float64 references: dense GEMM (3 operand layouts, TF32 and bf16)
and implicit-GEMM convolution fprop / dgrad / wgrad on the dpp.py layer shapes."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from distributeddataparallel_b200.ops import native

    native.load_library()
    yield native
    native.set_tensor_cores("auto")


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-12))


def _cl(t):
    return t.contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-3), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (256, 128, 512), (32, 16, 512), (1000, 200, 328), (4096, 1024, 1024),
                                   (16384, 2048, 512)])      # the last one takes the 128x256 tile path
def test_dense_gemm_three_layouts(nat, dtype, tol, m, n, k):
    torch.manual_seed(0)
    a = torch.randn(m, k, device="cuda").to(dtype)
    b_nk = torch.randn(n, k, device="cuda").to(dtype)
    ref = a.double() @ b_nk.double().t()
    out = nat.gemm_tc(a, b_nk, 0)
    assert _rel(out, ref) < tol, ("NT", _rel(out, ref))
    b_kn = b_nk.t().contiguous()
    out = nat.gemm_tc(a, b_kn, 1)
    assert _rel(out, ref) < tol, ("NN", _rel(out, ref))
    a_km = a.t().contiguous()
    out = nat.gemm_tc(a_km, b_kn, 2)
    assert _rel(out, ref) < tol, ("TN", _rel(out, ref))
    assert nat.tc_error_word() == 0


def test_dense_gemm_bias_and_accumulate(nat):
    a = torch.randn(64, 256, device="cuda")
    b = torch.randn(48, 256, device="cuda")
    bias = torch.randn(48, device="cuda")
    out = nat.gemm_tc(a, b, 0, bias=bias)
    ref = a.double() @ b.double().t() + bias.double()
    assert _rel(out, ref) < 2e-3
    out2 = nat.gemm_tc(a, b, 0, out=out.clone(), accumulate=True)
    assert _rel(out2, 2 * ref - bias.double()) < 2e-3


CONV_SHAPES = [
    (32, 64, 8, 64, 3, 1, 1), (32, 64, 8, 128, 3, 2, 1), (32, 128, 4, 128, 3, 1, 1), (32, 64, 8, 128, 1, 2, 0),
    (32, 128, 4, 256, 3, 2, 1), (32, 256, 2, 256, 3, 1, 1), (32, 256, 2, 512, 3, 2, 1), (32, 512, 1, 512, 3, 1, 1),
    (32, 256, 2, 512, 1, 2, 0), (10, 64, 8, 64, 3, 1, 1), (10, 512, 1, 512, 3, 1, 1), (8, 64, 56, 64, 3, 1, 1),
    (8, 256, 14, 1024, 1, 1, 0),       # pointwise, stride 1: dispatched as a dense (TMA) GEMM
    (32, 512, 1, 2048, 1, 1, 0),       # pointwise, tiny pixel count: swapped dense GEMM
    (16, 64, 56, 256, 1, 1, 0),
]


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_tc_matches_fp64(nat, shape):
    n, cin, h, cout, r, stride, pad = shape
    torch.manual_seed(1)
    nat.set_tensor_cores(True)
    before = nat.tc_launches
    x = _cl(torch.randn(n, cin, h, h, device="cuda")).requires_grad_(True)
    w = _cl(torch.randn(cout, cin, r, r, device="cuda") * 0.1).requires_grad_(True)
    stats = torch.zeros(2 * cout, device="cuda")
    y = nat.conv2d(x, w, stride, pad, stats)
    xd, wd = x.detach().double().requires_grad_(True), w.detach().double().requires_grad_(True)
    ref = F.conv2d(xd, wd, None, stride, pad)
    assert _rel(y, ref) < 2e-3, _rel(y, ref)
    assert torch.allclose(stats[:cout].double(), ref.sum(dim=(0, 2, 3)), rtol=2e-2, atol=0.5)
    assert torch.allclose(stats[cout:].double(), (ref * ref).sum(dim=(0, 2, 3)), rtol=2e-2, atol=0.5)
    dy = _cl(torch.randn_like(y))
    gx, gw = torch.autograd.grad(y, (x, w), dy)
    rx, rw = torch.autograd.grad(ref, (xd, wd), dy.double())
    assert _rel(gx, rx) < 2e-3 and _rel(gw, rw) < 2e-3, (_rel(gx, rx), _rel(gw, rw))
    assert nat.tc_launches - before == 3, "all three passes must have run on the tcgen05 kernel"
    assert nat.tc_error_word() == 0


def test_rgb_stem_runs_on_tensor_cores_with_a_zero_fourth_channel(nat):
    """C = 3 has no 16-byte channel vectors: the stem is run as a 4-channel convolution (zero channel added)
    when its input needs no gradient; with an input gradient requested it stays on the FFMA kernels."""
    nat.set_tensor_cores(True)
    x = _cl(torch.randn(8, 3, 32, 32, device="cuda"))
    w = _cl(torch.randn(64, 3, 7, 7, device="cuda")).requires_grad_(True)
    tc0, ff0 = nat.tc_launches, nat.ffma_launches
    y = nat.conv2d(x, w, 2, 3)
    dy = torch.randn_like(y)
    y.backward(dy)
    assert nat.tc_launches == tc0 + 2 and nat.ffma_launches == ff0          # fprop + wgrad on tcgen05
    xd, wd = x.double(), w.detach().double().requires_grad_(True)
    yd = F.conv2d(xd, wd, None, 2, 3)
    yd.backward(dy.double())
    assert _rel(y, yd) < 2e-3 and _rel(w.grad, wd.grad) < 2e-3
    assert w.grad.shape == w.shape
    xg = x.clone().requires_grad_(True)
    ff1 = nat.ffma_launches
    y2 = nat.conv2d(xg, w, 2, 3)
    assert nat.ffma_launches == ff1 + 1
    assert _rel(y2, yd) < 1e-5


@pytest.mark.parametrize("c,k,r,pad", [(64, 128, 3, 1), (64, 128, 1, 0)])
def test_stride2_dgrad_by_parity_classes(nat, c, k, r, pad):
    """Stride-2 dgrad of a large image runs as four stride-1 problems (one per input-pixel parity) through the
    TMA-fed forward kernels and is interleaved into dX; it must agree with the gather kernel and with fp64."""
    import ctypes

    lib = nat.load_library()
    nat.set_tensor_cores(True)
    torch.manual_seed(0)
    x = _cl(torch.randn(8, c, 56, 56, device="cuda")).requires_grad_(True)        # 25 088 input pixels >= 16 384
    w = _cl(torch.randn(k, c, r, r, device="cuda") * 0.05).requires_grad_(True)
    y = nat.conv2d(x, w, 2, pad)
    dy = _cl(torch.randn_like(y))
    got = {}
    try:
        for flag in (1, 0):
            lib.b200_gemm_tc_use_s2_dgrad(ctypes.c_int(flag))
            (got[flag],) = torch.autograd.grad(nat.conv2d(x, w, 2, pad), x, dy)
    finally:
        lib.b200_gemm_tc_use_s2_dgrad(ctypes.c_int(1))
    ref = torch.autograd.grad(F.conv2d(x.double(), w.double(), None, 2, pad), x, dy.double())[0]
    assert _rel(got[1], ref) < 2e-3 and _rel(got[0], ref) < 2e-3
    assert _rel(got[1], got[0]) < 2e-3
    assert nat.tc_error_word() == 0
