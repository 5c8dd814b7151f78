"""Print what the convolution auto-tuner picked for every shape of the dpp.py model (0 = FFMA kernel;
otherwise split-K bound of the tcgen05 kernel, +100 = splits reduced through a cluster's shared memory)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import distributeddataparallel_b200 as d
from distributeddataparallel_b200.ops import native

torch.manual_seed(0)
model = d.DDP(d.SimpleCNN().cuda(), device_ids=[0])
opt = d.SGD(model.parameters(), lr=0.01)
step = d.TrainStep(model, d.CrossEntropyLoss(), opt, cuda_graph=False)
x = d.ops.preprocess_u8(torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, device="cuda"))
y = torch.randint(0, 10, (32,), device="cuda")
for _ in range(3):
    step(x, y)
torch.cuda.synchronize()
for key, v in sorted(native.conv_choices().items(), key=lambda kv: (kv[0][0], kv[0][1:])):
    name = key[0].replace("b200_conv2d_", "")
    n, h, w, c, k, r, s, stride, pad = key[1:]
    print(f"{name:6s} N{n} {h}x{w} C{c} K{k} R{r} s{stride} p{pad} -> {v}")
