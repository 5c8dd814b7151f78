import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.getcwd())
rank = int(os.environ["RANK"]); ws = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
import distributeddataparallel_b200 as d
from distributeddataparallel_b200.parallel.p2p import P2PComm
d.setup(rank, ws, backend="nccl")
dev = torch.device("cuda", rank)
comm = P2PComm(None, dev)
comm.arena and None
comm.reserve(8 << 20)
comm.arena.set_timeout(5.0)
flat = comm.allocate(1 << 20, torch.float32, dev)
flat.fill_(float(rank + 1))
comm.launch(0, flat, True).wait()
torch.cuda.synchronize()
print(rank, comm.arena.provider, bool(comm.arena.multicast), float(flat[0]), float(flat[-1]), comm.arena.error_word(), flush=True)
os._exit(0)
