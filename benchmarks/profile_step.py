"""Run a few eager training steps of the dpp.py model with the profiler gate open around the
last ones (for `ncu --profile-from-start off`): gives one line per kernel launch of a step."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import distributeddataparallel_b200 as d

p = argparse.ArgumentParser()
p.add_argument("--steps", type=int, default=1)
p.add_argument("--warmup", type=int, default=4)
p.add_argument("--batch", type=int, default=32)
p.add_argument("--model", default="simplecnn")
a = p.parse_args()

torch.manual_seed(0)
cfg = d.TrainConfig(model=a.model)
model = d.DDP(d.build_model(cfg).cuda(), device_ids=[0])
opt = d.SGD(model.parameters(), lr=0.01)
step = d.TrainStep(model, d.CrossEntropyLoss(), opt, cuda_graph=False)
x = d.ops.preprocess_u8(torch.randint(0, 256, (a.batch, 32, 32, 3), dtype=torch.uint8, device="cuda"))
y = torch.randint(0, 10, (a.batch,), device="cuda")
for _ in range(a.warmup):
    step(x, y)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(a.steps):
    step(x, y)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
