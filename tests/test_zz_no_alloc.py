"""The allocation-free forward convolution, driven through the C ABI exactly as a foreign host
would drive it.

Who needs it: callers that may not allocate device memory at the moment they need the op -- a
thread working while another stream of the process is being captured into a CUDA graph, a
plug-in living inside a runtime that owns the memory budget, a loop that has frozen its
allocations.  ``b200_conv2d_fprop_noalloc`` takes inputs and outputs only; its split-K scratch
is a fixed array inside the library.

Checked here: on two calls in a row (the second on different data, so anything stale in the
shared scratch would show) its output equals that of ``b200_conv2d_fprop`` -- the entry point
that takes caller scratch -- within the tolerance the convolution tests use; the fused
statistics agree too; and a problem beyond the stated limit is refused with a readable message.

The file name sorts last on purpose: a ``-x`` run reaches it only after every other test ran.
It does not depend on them: it loads the library and creates its own tensors.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5          # same relative-norm bound as tests/test_kernels_gpu.py::test_conv_fprop_dgrad_wgrad

_vp, _i, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
_GEOM = [_i] * 9


def _lib():
    from distributeddataparallel_b200 import _build

    lib = ctypes.CDLL(str(_build.build_kernels()))
    lib.b200_conv2d_fprop.restype = _i
    lib.b200_conv2d_fprop.argtypes = [_vp, _vp, _vp, _vp, _vp] + _GEOM + [_vp, _sz, _vp, _i, _vp]
    lib.b200_conv2d_fprop_noalloc.restype = _i
    lib.b200_conv2d_fprop_noalloc.argtypes = [_vp, _vp, _vp, _vp, _vp] + _GEOM + [_vp]
    lib.b200_conv2d_fprop_noalloc_error.restype = ctypes.c_char_p
    lib.b200_conv2d_fprop_noalloc_scratch_bytes.restype = _sz
    return lib


def _rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-12))


# (N, H, C, K, R, stride, pad): deep split-K (M = 32, 32 splits), mid split-K, a 2-way split (128 tiles on 148
# SMs), and no split at all (256 tiles: the scratch array is not touched)
SHAPES = [(32, 1, 512, 512, 3, 1, 1), (32, 8, 64, 64, 3, 1, 1), (32, 32, 3, 64, 7, 2, 3), (64, 16, 16, 64, 3, 1, 1)]


@pytest.mark.parametrize("shape", SHAPES)
def test_noalloc_conv_matches_scratch_taking_entry_point(shape):
    lib = _lib()
    n, h, c, k, r, stride, pad = shape
    p = (h + 2 * pad - r) // stride + 1
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    workspace = torch.empty(8 << 20, dtype=torch.float32, device=dev)
    tickets = torch.zeros(4096, dtype=torch.int32, device=dev)
    g = torch.Generator(device="cpu").manual_seed(11)
    for call in range(2):
        x = torch.randn(n, h, h, c, generator=g).to(dev)                # NHWC
        w = (torch.randn(k, r, r, c, generator=g) * 0.1).to(dev)        # KRSC
        bias = torch.randn(k, generator=g).to(dev)
        y_ref = torch.empty(n, p, p, k, device=dev)
        y_new = torch.full_like(y_ref, float("nan"))
        s_ref = torch.zeros(2 * k, device=dev)
        s_new = torch.zeros(2 * k, device=dev)
        geom = (n, h, h, c, k, r, r, stride, pad)
        rc = lib.b200_conv2d_fprop(x.data_ptr(), w.data_ptr(), bias.data_ptr(), y_ref.data_ptr(), s_ref.data_ptr(),
                                   *geom, workspace.data_ptr(), workspace.numel() * 4, tickets.data_ptr(),
                                   tickets.numel(), stream)
        assert rc == 0
        rc = lib.b200_conv2d_fprop_noalloc(x.data_ptr(), w.data_ptr(), bias.data_ptr(), y_new.data_ptr(),
                                           s_new.data_ptr(), *geom, stream)
        assert rc == 0, lib.b200_conv2d_fprop_noalloc_error().decode()
        torch.cuda.synchronize()
        assert torch.isfinite(y_new).all()
        assert _rel(y_new, y_ref) < TOL, (call, _rel(y_new, y_ref))
        assert torch.allclose(s_new, s_ref, rtol=1e-4, atol=1e-3), call
        # and both agree with a float64 convolution of the same data
        ref64 = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2).double(), w.permute(0, 3, 1, 2).double(),
                                           bias.double(), stride, pad).permute(0, 2, 3, 1)
        assert _rel(y_new.double(), ref64) < TOL
    assert int(tickets.abs().sum()) == 0


def test_noalloc_conv_refuses_oversized_problem_with_message():
    lib = _lib()
    assert lib.b200_conv2d_fprop_noalloc_scratch_bytes() == (4 << 20) + 4096
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    # 64 x 64 x 64 output pixels x 64 channels = 4096 tiles > 1024: refused before touching memory
    dummy = torch.zeros(16, device=dev)
    rc = lib.b200_conv2d_fprop_noalloc(dummy.data_ptr(), dummy.data_ptr(), None, dummy.data_ptr(), None,
                                       64, 64, 64, 4, 64, 1, 1, 1, 0, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.b200_conv2d_fprop_noalloc_error().decode()
    assert "4096 tiles" in msg and "1024" in msg and "b200_conv2d_fprop" in msg
    torch.cuda.synchronize()
