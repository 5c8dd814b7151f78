"""Per-kernel timeline of the dpp.py training step under multi-GPU DDP (run with torchrun); rank 0 prints.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 benchmarks/timeline_ddp.py
"""
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

import distributeddataparallel_b200 as d

rank, ws = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
d.setup(rank, ws)
torch.manual_seed(0)
model = d.DDP(d.SimpleCNN().cuda(), device_ids=[rank])
opt = d.SGD(model.parameters(), lr=0.01)
step = d.TrainStep(model, d.CrossEntropyLoss(), opt, cuda_graph=True)
x = d.ops.preprocess_u8(torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, device="cuda"))
y = torch.randint(0, 10, (32,), device="cuda")
for _ in range(12):
    step(x, y)
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    step(x, y)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        step(x, y)
    torch.cuda.synchronize()
if rank == 0:
    ev = sorted((e for e in prof.events() if e.device_type.name == "CUDA"), key=lambda e: e.time_range.start)
    agg = collections.OrderedDict()
    for e in ev:
        a = agg.setdefault(e.name[:80], [0, 0.0])
        a[0] += 1
        a[1] += e.device_time
    tot = sum(v[1] for v in agg.values())
    print(f"world {ws}: {ms * 1e3:.1f} us/step; kernel time per step {tot / 5:.1f} us over {len(ev) // 5} kernels; comm = {model.comm.name}")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
        print(f"{t / 5:8.1f} us/step  n={n // 5:3d}  avg={t / n:6.2f}  {k}")
    last = ev[len(ev) * 4 // 5:]
    t0 = last[0].time_range.start
    out = os.environ.get("B200DDP_SEQ_OUT")
    if out:
        with open(out, "w") as f:
            for e in last:
                f.write(f"{e.time_range.start - t0:9.1f} {e.device_time:7.1f} {e.name[:90]}\n")
d.cleanup()
