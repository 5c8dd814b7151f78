"""Warm, in-graph per-kernel timeline of the dpp.py training step (torch.profiler / CUPTI on graph replays)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
import distributeddataparallel_b200 as d

torch.manual_seed(0)
model = d.DDP(d.SimpleCNN().cuda(), device_ids=[0])
opt = d.SGD(model.parameters(), lr=0.01)
step = d.TrainStep(model, d.CrossEntropyLoss(), opt, cuda_graph=True)
x = d.ops.preprocess_u8(torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, device="cuda"))
y = torch.randint(0, 10, (32,), device="cuda")
for _ in range(12):
    step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(5):
        step(x, y)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type.name == "CUDA"]
import collections
agg = collections.OrderedDict()
for e in ev:
    a = agg.setdefault(e.name[:80], [0, 0.0])
    a[0] += 1; a[1] += e.device_time if hasattr(e, "device_time") else e.cuda_time
tot = sum(v[1] for v in agg.values())
print(f"total kernel time per step: {tot/5:.1f} us over {sum(v[0] for v in agg.values())//5} kernels")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
    print(f"{t/5:8.1f} us/step  n={n//5:3d}  avg={t/n:6.2f}  {k}")
ts = sorted((e.time_range.start, e.time_range.end) for e in ev)
span = (ts[-1][1] - ts[0][0]) / 5
print(f"span per step ~{span:.1f} us")
