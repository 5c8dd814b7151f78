"""Run one convolution pass of the dpp.py model on the tcgen05 kernel (for ncu captures / quick timing).

    python benchmarks/conv_probe.py [N C H W K R stride pad] [--pass fprop|dgrad|wgrad]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributeddataparallel_b200.ops import native

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n, c, h, w, k, r, stride, pad = (int(v) for v in args) if len(args) == 8 else (32, 64, 8, 8, 64, 3, 1, 1)
which = "fprop"
if "--pass" in sys.argv:
    which = sys.argv[sys.argv.index("--pass") + 1]
native.set_tensor_cores(True)
x = torch.randn(n, c, h, w, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
wt = torch.randn(k, c, r, r, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
y = native.conv2d(x, wt, stride, pad)
dy = torch.randn_like(y)


def run():
    if which == "fprop":
        with torch.no_grad():
            native.conv2d(x, wt, stride, pad)
    else:
        x.grad = wt.grad = None
        yy = native.conv2d(x, wt, stride, pad)
        torch.autograd.grad(yy, x if which == "dgrad" else wt, dy)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(20):
        run()
g.replay()
torch.cuda.synchronize()
e0.record()
g.replay()
e1.record()
torch.cuda.synchronize()
print(f"{which} N{n} C{c} {h}x{w} K{k} R{r} s{stride} p{pad}: {e0.elapsed_time(e1) / 20 * 1e3:.2f} us per call (graph, back to back)")
