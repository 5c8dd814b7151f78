"""Top-level worker functions for mp.spawn based tests (must be importable by name)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist


def _init(rank, ws, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(2)
    import distributeddataparallel_b200 as d

    d.setup(rank, ws, backend="gloo")
    return d


class TinyNet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(8, 16)
        self.bn = torch.nn.BatchNorm1d(16)
        self.b = torch.nn.Linear(16, 4)
        self.unused = torch.nn.Linear(3, 3)

    def forward(self, x, use_unused=False):
        y = self.b(torch.relu(self.bn(self.a(x))))
        if use_unused:
            y = y + self.unused(x[:, :3]).sum(dim=1, keepdim=True)
        return y


def ddp_parity_worker(rank, ws, port, out_dir, steps, channels_last=False, tail_splits=None):
    """Our DDP vs stock DDP on the reference model: losses, grads, params, collective schedule."""
    d = _init(rank, ws, port)
    from torch.nn.parallel import DistributedDataParallel as StockDDP

    torch.manual_seed(100 + rank)            # different init per rank: ctor broadcast must fix it
    ours = d.SimpleCNN(channels_last=channels_last)
    mine = d.DDP(ours, comm="c10d", tail_splits=tail_splits)
    stock_model = d.SimpleCNN(channels_last=False)
    d.ops.set_native(False)
    stock_model.load_state_dict(ours.state_dict())
    stock = StockDDP(stock_model)
    opt_a = d.SGD(mine.parameters(), lr=0.01)
    opt_b = torch.optim.SGD(stock.parameters(), lr=0.01)
    g = torch.Generator().manual_seed(7 + rank)
    rec = {"loss_a": [], "loss_b": [], "buckets": [], "coll": []}
    p0 = torch.cat([p.detach().reshape(-1)[:64] for p in ours.parameters()])
    gathered = [torch.zeros_like(p0) for _ in range(ws)]
    dist.all_gather(gathered, p0)
    rec["ctor_equal"] = all(torch.equal(gathered[0], t) for t in gathered)
    for it in range(steps):
        x = torch.randn(8, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (8,), generator=g)
        before = mine._comm.launches
        opt_a.zero_grad()
        la = d.CrossEntropyLoss()(mine(x), y)
        la.backward()
        rec["buckets"].append(mine._comm.launches - before)
        opt_b.zero_grad()
        lb = torch.nn.functional.cross_entropy(stock(x), y)
        lb.backward()
        ga = torch.cat([p.grad.reshape(-1) for p in mine.parameters()])
        gb = torch.cat([p.grad.reshape(-1) for p in stock.parameters()])
        rec.setdefault("grad_err", []).append(float((ga - gb).norm() / gb.norm()))
        gl = [torch.zeros_like(ga) for _ in range(ws)]
        dist.all_gather(gl, ga)
        rec.setdefault("grads_identical", []).append(all(torch.equal(gl[0], t) for t in gl))
        opt_a.step()
        opt_b.step()
        rec["loss_a"].append(float(la))
        rec["loss_b"].append(float(lb))
    log = mine._get_ddp_logging_data()
    rec["log"] = {k: v for k, v in log.items() if isinstance(v, (int, float, str, bool, list, dict))}
    rec["keys"] = list(mine.state_dict().keys())
    pa = torch.cat([p.detach().reshape(-1) for p in mine.parameters()])
    pb = torch.cat([p.detach().reshape(-1) for p in stock.parameters()])
    rec["param_err"] = float((pa - pb).norm() / pb.norm())
    torch.save(rec, os.path.join(out_dir, f"rank{rank}.pt"))
    d.cleanup()


def features_worker(rank, ws, port, out_dir):
    """no_sync, comm hooks, find_unused_parameters, mismatch detection."""
    d = _init(rank, ws, port)
    from distributeddataparallel_b200.parallel import comm as C

    rec = {}
    torch.manual_seed(0)
    net = TinyNet()
    net.unused.requires_grad_(False)
    ddp = d.DDP(net, comm="c10d")
    g = torch.Generator().manual_seed(rank)
    xs = [torch.randn(6, 8, generator=g) for _ in range(4)]
    # --- no_sync: two local accumulations then one synced step == mean of summed grads
    with ddp.no_sync():
        ddp(xs[0]).sum().backward()
    before = ddp._comm.launches
    ddp(xs[1]).sum().backward()
    rec["nosync_launches"] = ddp._comm.launches - before
    got = net.a.weight.grad.clone()
    ref_net = TinyNet()
    ref_net.load_state_dict(net.state_dict())
    ref_net.unused.requires_grad_(False)
    # BN running stats advanced in `net`; grads only depend on batch stats so reuse is fine
    ref_net(xs[0]).sum().backward()
    ref_net(xs[1]).sum().backward()
    local = ref_net.a.weight.grad.clone()
    dist.all_reduce(local)
    rec["nosync_err"] = float((got - local / ws).abs().max())
    # --- comm hook (bf16 compression)
    net2 = TinyNet()
    net2.unused.requires_grad_(False)
    ddp2 = d.DDP(net2, comm="c10d")
    ddp2.register_comm_hook(None, C.bf16_compress_hook)
    ddp2(xs[2]).sum().backward()
    gl = [torch.zeros_like(net2.a.weight.grad) for _ in range(ws)]
    dist.all_gather(gl, net2.a.weight.grad)
    rec["hook_identical"] = all(torch.equal(gl[0], t) for t in gl)
    rec["hook_name"] = ddp2._get_ddp_logging_data()["comm_hook"]
    # --- find_unused_parameters
    net3 = TinyNet()
    ddp3 = d.DDP(net3, comm="c10d", find_unused_parameters=True)
    ddp3(xs[3], use_unused=False).sum().backward()
    ddp3(xs[3], use_unused=(rank == 0)).sum().backward()
    rec["unused_grad_zero"] = bool(net3.unused.weight.grad is not None)
    gl = [torch.zeros_like(net3.unused.weight.grad) for _ in range(ws)]
    dist.all_gather(gl, net3.unused.weight.grad)
    rec["unused_identical"] = all(torch.equal(gl[0], t) for t in gl)
    # --- unused parameter without the flag -> clear error on the next iteration
    net4 = TinyNet()
    ddp4 = d.DDP(net4, comm="c10d")
    try:
        ddp4(xs[3]).sum().backward()
        rec["unfinished_error"] = ""
    except RuntimeError as e:
        rec["unfinished_error"] = str(e)
    # --- shape mismatch across ranks is detected in the constructor
    bad = torch.nn.Linear(4, 4 + rank)
    try:
        d.DDP(bad, comm="c10d")
        rec["mismatch_error"] = ""
    except RuntimeError as e:
        rec["mismatch_error"] = str(e)
    torch.save(rec, os.path.join(out_dir, f"rank{rank}.pt"))
    d.cleanup()


def train_worker(rank, ws, port, out_dir, ckpt, resume):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(2)
    import distributeddataparallel_b200 as d

    cfg = d.TrainConfig(backend="gloo", device="cpu", dataset_size=256, batch_size=16, max_steps=6,
                        log_interval=2, checkpoint_path=ckpt, resume=resume, master_port=port, comm="c10d")
    stats = d.train(rank, ws, epochs=2, cfg=cfg)
    torch.save(stats, os.path.join(out_dir, f"train{rank}.pt"))


def failing_worker(rank, ws, port):
    d = _init(rank, ws, port)
    if rank == 1:
        raise ValueError("injected failure on rank 1")
    dist.barrier()          # rank 0 blocks here until torn down by the launcher
    d.cleanup()


def join_worker(rank, ws, port, out_dir, divide_by_world=True):
    """Uneven inputs: rank r trains 2 + 3*r iterations inside join(); ours vs stock DDP's join()."""
    d = _init(rank, ws, port)
    from torch.nn.parallel import DistributedDataParallel as StockDDP

    d.ops.set_native(False)
    torch.manual_seed(5)
    a, b = TinyNet(), TinyNet()
    b.load_state_dict(a.state_dict())
    mine = d.DDP(a, comm="c10d", bucket_cap_mb=0.0005)       # several buckets + a rebuild inside the join
    stock = StockDDP(b, bucket_cap_mb=0.0005)
    opt_a = torch.optim.SGD(mine.parameters(), lr=0.05)
    opt_b = torch.optim.SGD(stock.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(11 + rank)
    batches = [(torch.randn(6, 8, generator=g), torch.randn(6, 4, generator=g)) for _ in range(2 + 3 * rank)]
    for model, opt in ((mine, opt_a), (stock, opt_b)):
        with model.join(divide_by_initial_world_size=divide_by_world):
            for x, y in batches:
                opt.zero_grad()
                loss = ((model(x)[:, :4] - y) ** 2).mean() + 0.0 * sum(p.sum() for p in model.module.unused.parameters())
                loss.backward()
                opt.step()
    pa = torch.cat([p.detach().reshape(-1) for p in mine.parameters()] + [t.detach().float().reshape(-1) for t in mine.buffers()])
    pb = torch.cat([p.detach().reshape(-1) for p in stock.parameters()] + [t.detach().float().reshape(-1) for t in stock.buffers()])
    gathered = [torch.zeros_like(pa) for _ in range(ws)]
    dist.all_gather(gathered, pa)
    rec = {"identical_across_ranks": all(torch.equal(gathered[0], t) for t in gathered),
           "err_vs_stock": float((pa - pb).norm() / pb.norm()),
           "param_err": float(max((x.detach() - y.detach()).abs().max() for x, y in zip(mine.parameters(), stock.parameters()))),
           "control": mine._get_ddp_logging_data()["collectives"]["allreduce_control"],
           "iters": len(batches), "divisor_after": mine.comm.divisor}
    torch.save(rec, os.path.join(out_dir, f"join{rank}.pt"))
    d.cleanup()


def join_throw_worker(rank, ws, port, out_dir):
    """join(throw_on_early_termination=True): rank 0 runs out after 2 iterations -> every rank raises."""
    d = _init(rank, ws, port)
    d.ops.set_native(False)
    torch.manual_seed(5)
    mine = d.DDP(TinyNet(), comm="c10d")
    opt = torch.optim.SGD(mine.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(11 + rank)
    done, msg = 0, ""
    try:
        with mine.join(throw_on_early_termination=True):
            for _ in range(2 + 3 * rank):
                x, y = torch.randn(6, 8, generator=g), torch.randn(6, 4, generator=g)
                opt.zero_grad()
                (((mine(x) - y) ** 2).mean() + 0.0 * sum(p.sum() for p in mine.module.unused.parameters())).backward()
                opt.step()
                done += 1
    except RuntimeError as exc:
        msg = str(exc)
    torch.save({"done": done, "msg": msg}, os.path.join(out_dir, f"throw{rank}.pt"))
    d.cleanup()


def ignore_worker(rank, ws, port, out_dir):
    """module._ddp_params_and_buffers_to_ignore: the listed parameter / buffer is neither broadcast nor reduced."""
    d = _init(rank, ws, port)
    d.ops.set_native(False)
    torch.manual_seed(100 + rank)
    net = TinyNet()
    d.DDP._set_params_and_buffers_to_ignore_for_model(net, ["unused.weight", "bn.running_mean"])
    before = net.unused.weight.detach().clone()
    mine = d.DDP(net, comm="c10d")
    kept_local = torch.equal(net.unused.weight.detach(), before)
    x = torch.randn(6, 8, generator=torch.Generator().manual_seed(rank))
    mine(x, use_unused=True).sum().backward()
    gl = [torch.zeros_like(net.unused.weight.grad) for _ in range(ws)]
    dist.all_gather(gl, net.unused.weight.grad)
    al = [torch.zeros_like(net.a.weight.grad) for _ in range(ws)]
    dist.all_gather(al, net.a.weight.grad)
    rm = [torch.zeros_like(net.bn.running_mean) for _ in range(ws)]
    dist.all_gather(rm, net.bn.running_mean)
    mine(x, use_unused=True)                      # a second forward would have re-broadcast rank 0's buffers
    rv = [torch.zeros_like(net.bn.running_var) for _ in range(ws)]
    dist.all_gather(rv, net.bn.running_var)
    torch.save({"kept_local": kept_local, "ignored_grad_differs": not torch.equal(gl[0], gl[1]),
                "other_grad_equal": torch.equal(al[0], al[1]), "ignored_buffer_differs": not torch.equal(rm[0], rm[1]),
                "names": mine._param_names, "will_sync": mine.will_sync_module_buffers()},
               os.path.join(out_dir, f"ign{rank}.pt"))
    d.cleanup()


def lm_train_worker(rank, ws, port, out_dir):
    """``--model gpt2-tiny`` through train(): token dataset, fused-loss model, DDP, SGD."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(2)
    import distributeddataparallel_b200 as d

    cfg = d.TrainConfig(backend="gloo", device="cpu", model="gpt2-tiny", dataset_size=64, seq_len=32, batch_size=4,
                        max_steps=4, log_interval=1, master_port=port, comm="c10d")
    stats = d.train(rank, ws, epochs=1, cfg=cfg)
    torch.save(stats, os.path.join(out_dir, f"lm{rank}.pt"))
