"""Driver-specified extra workloads (BASELINE.json configs 3 and 4): ResNet-50 (ImageNet-shaped
batches) and GPT-2 medium under data parallelism, ours vs the stock PyTorch stack.

    [torchrun --nproc-per-node N] python benchmarks/bench_models.py --model resnet50|gpt2-medium \
        --impl ours|stock [--batch B] [--steps K] [--warmup W] [--dtype bf16|tf32]

Prints one JSON line (samples/s aggregate over ranks, device-timed, max over ranks).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

p = argparse.ArgumentParser()
p.add_argument("--model", default="gpt2-medium")
p.add_argument("--impl", default="ours", choices=["ours", "stock"])
p.add_argument("--batch", type=int, default=None)
p.add_argument("--steps", type=int, default=10)
p.add_argument("--warmup", type=int, default=4)
p.add_argument("--dtype", default=None)
p.add_argument("--seq-len", type=int, default=1024)
p.add_argument("--no-graph", action="store_true")
a = p.parse_args()

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29577")
saved = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
torch.manual_seed(0)
is_gpt = a.model.startswith("gpt2")
batch = a.batch or (8 if is_gpt else 32)
dtype = a.dtype or ("bf16" if is_gpt else "tf32")

if a.impl == "ours":
    import distributeddataparallel_b200 as d

    if is_gpt:
        from distributeddataparallel_b200.models.gpt2 import build_gpt2

        model = build_gpt2(a.model, a.seq_len).to(dev, torch.bfloat16 if dtype == "bf16" else torch.float32)
        x = torch.randint(0, 50257, (batch, a.seq_len), device=dev)
        y = torch.roll(x, -1, dims=1)
        crit = None
    else:
        model = d.resnet50(1000).to(dev)
        x = torch.randn(batch, 3, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 1000, (batch,), device=dev)
        crit = d.CrossEntropyLoss()
    ddp = d.DDP(model, device_ids=[local])
    opt = d.SGD(ddp.parameters(), lr=0.01)
    step = d.TrainStep(ddp, crit, opt, cuda_graph=not a.no_graph)
    run = lambda: step(x, y)
    comm = ddp.comm.name
else:
    from torch.nn.parallel import DistributedDataParallel as StockDDP

    if is_gpt:
        from transformers import GPT2Config, GPT2LMHeadModel

        cfg = GPT2Config(n_layer=24, n_head=16, n_embd=1024, n_positions=a.seq_len, vocab_size=50257,
                         attn_pdrop=0.0, embd_pdrop=0.0, resid_pdrop=0.0) if a.model == "gpt2-medium" else GPT2Config()
        cfg._attn_implementation = "sdpa"
        model = GPT2LMHeadModel(cfg).to(dev, torch.bfloat16 if dtype == "bf16" else torch.float32)
        x = torch.randint(0, 50257, (batch, a.seq_len), device=dev)
        fwd = lambda: model_ddp(input_ids=x, labels=x).loss
    else:
        import torchvision

        model = torchvision.models.resnet50(weights=None).to(dev).to(memory_format=torch.channels_last)
        x = torch.randn(batch, 3, 224, 224, device=dev).contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 1000, (batch,), device=dev)
        if dtype == "bf16":
            def fwd():
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    return torch.nn.functional.cross_entropy(model_ddp(x), y)
        else:
            fwd = lambda: torch.nn.functional.cross_entropy(model_ddp(x), y)
    model_ddp = StockDDP(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.SGD(model.parameters(), lr=0.01)

    def run():
        opt.zero_grad()
        loss = fwd()
        loss.backward()
        opt.step()
        return loss.detach()
    comm = "nccl (torch DDP)"

for _ in range(a.warmup + 4):
    loss = run()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    loss = run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
if world > 1:
    t = torch.tensor([ms], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
out = {"model": a.model, "impl": a.impl, "dtype": dtype, "n_gpus": world, "per_gpu_batch": batch,
       "samples_per_s": a.steps * batch * world / (ms / 1e3), "ms_per_step": ms / a.steps, "loss": float(loss),
       "comm": comm, "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30}
if is_gpt:
    out["tokens_per_s"] = out["samples_per_s"] * a.seq_len
if rank == 0:
    saved.write(json.dumps(out) + "\n")
    saved.flush()
if world > 1:
    dist.destroy_process_group()
