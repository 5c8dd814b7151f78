"""Time the BatchNorm kernels (fwd apply, bwd reduce, bwd apply) on one ResNet-50 activation shape and report the
achieved HBM bandwidth.   python benchmarks/bn_probe.py [N C H W] [--residual]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributeddataparallel_b200.ops import native

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n, c, h, w = (int(v) for v in args) if len(args) == 4 else (32, 256, 56, 56)
res = "--residual" in sys.argv
x = torch.randn(n, c, h, w, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
r = torch.randn_like(x) if res else None
bn = torch.nn.BatchNorm2d(c).cuda()


def fwd():
    return native.batch_norm_act(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked, r, True,
                                 True, 0.1, 1e-5, native.bn_workspace(bn, c, x.device), False)


y = fwd()
dy = torch.randn_like(y)
big = torch.empty(200 << 20, dtype=torch.uint8, device="cuda")
from torch.profiler import ProfilerActivity, profile

for _ in range(2):
    x.grad = None
    fwd().backward(dy)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        big.zero_()
        x.grad = None
        fwd().backward(dy)
    torch.cuda.synchronize()
agg = {}
for e in prof.events():
    if e.device_type.name == "CUDA" and "bn_" in e.name:
        a = agg.setdefault(e.name.split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += e.device_time
nbytes = x.numel() * 4
traffic = {"bn_stats": 1, "bn_fwd_apply": 2 + int(res), "bn_bwd_reduce": 3, "bn_bwd_apply": 4 + int(res)}
for k, (cnt, t) in agg.items():
    key = next((kk for kk in traffic if kk in k), None)
    us = t / cnt
    bw = traffic[key] * nbytes / us / 1e6 if key else 0
    print(f"{k:40s} {us:8.1f} us   ~{bw:6.2f} TB/s of {traffic.get(key, 0)} x {nbytes / 1e6:.0f} MB")
