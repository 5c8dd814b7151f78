"""Per-kernel time breakdown of one GPT-2 medium training step on this repo's kernels (torch.profiler)."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
import distributeddataparallel_b200 as d
from distributeddataparallel_b200.models.gpt2 import build_gpt2

model_name = sys.argv[1] if len(sys.argv) > 1 else "gpt2-medium"
torch.manual_seed(0)
if model_name.startswith("gpt2"):
    model = build_gpt2(model_name, 1024).to("cuda", torch.bfloat16)
    x = torch.randint(0, 50257, (8, 1024), device="cuda"); y = torch.roll(x, -1, 1); crit = None
else:
    model = d.resnet50(1000).cuda()
    x = torch.randn(32, 3, 224, 224, device="cuda").contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (32,), device="cuda"); crit = d.CrossEntropyLoss()
ddp = d.DDP(model, device_ids=[0])
opt = d.SGD(ddp.parameters(), lr=0.01)
step = d.TrainStep(ddp, crit, opt, cuda_graph=False)
for _ in range(4):
    step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step(x, y)
    torch.cuda.synchronize()
agg = collections.OrderedDict()
for e in prof.events():
    if e.device_type.name != "CUDA":
        continue
    a = agg.setdefault(e.name[:95], [0, 0.0])
    a[0] += 1; a[1] += e.device_time
if os.environ.get("B200DDP_SEQ_OUT"):
    # every kernel of the LAST profiled step, in launch order, with its duration
    evs = sorted((e for e in prof.events() if e.device_type.name == "CUDA"), key=lambda e: e.time_range.start)
    evs = evs[len(evs) // 2:]
    with open(os.environ["B200DDP_SEQ_OUT"], "w") as f:
        t0 = evs[0].time_range.start
        for e in evs:
            f.write(f"{e.time_range.start - t0:10.1f} {e.device_time:8.1f} {e.name[:110]}\n")
tot = sum(v[1] for v in agg.values())
print(f"kernel time per step {tot/2/1e3:.2f} ms")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{t/2/1e3:8.2f} ms {100*t/tot:5.1f}%  n={n//2:4d} avg={t/n:8.1f}us  {k}")
