import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from distributeddataparallel_b200.ops import native
lib = native.load_library()
native.set_tensor_cores(True)
def cl(t): return t.contiguous(memory_format=torch.channels_last)
def rel(a, b): return float((a.double() - b.double()).norm() / b.double().norm())
ok = True
for (n, c, h, k, r, p) in [(32, 128, 56, 128, 3, 1), (32, 256, 56, 512, 1, 0), (32, 64, 56, 128, 3, 1)]:
    torch.manual_seed(0)
    x = cl(torch.randn(n, c, h, h, device="cuda")).requires_grad_(True)
    w = cl(torch.randn(k, c, r, r, device="cuda") * 0.05).requires_grad_(True)
    res = {}
    for flag in (0, 1):
        lib.b200_gemm_tc_use_s2_dgrad(ctypes.c_int(flag))
        y = native.conv2d(x, w, 2, p)
        dy = cl(torch.randn_like(y)) if flag == 0 else dy
        (gx,) = torch.autograd.grad(y, x, dy)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        y = native.conv2d(x, w, 2, p)
        e0.record(); (gx,) = torch.autograd.grad(y, x, dy); e1.record(); torch.cuda.synchronize()
        res[flag] = (gx.clone(), e0.elapsed_time(e1) * 1e3)
    lib.b200_gemm_tc_use_s2_dgrad(ctypes.c_int(0))
    ref = torch.autograd.grad(F.conv2d(x.double(), w.double(), None, 2, p), x, dy.double())[0]
    e_old, e_new = rel(res[0][0], ref), rel(res[1][0], ref)
    print(f"N{n} C{c} {h}x{h} K{k} R{r}: gather {res[0][1]:.1f} us err {e_old:.2e} | parity {res[1][1]:.1f} us err {e_new:.2e}  tc_error={native.tc_error_word()}")
    ok &= e_new < 3e-3
print("S2 OK" if ok else "S2 FAIL")
