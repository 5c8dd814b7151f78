"""Symmetric arena + the P2P (NVLink) communication backend.

This is the product data plane: gradient buckets, module buffers and broadcast staging live in
a per-rank *symmetric arena* -- one ``cudaMalloc`` block whose CUDA-IPC handle every peer has
opened, so that an object sits at the same offset in every rank's arena and each rank can
``ld``/``st`` its peers' copies directly over NVLink 5 / NVSwitch.  The collectives are the
kernels in ``csrc/kernels/comm.cu`` (barrier, one-shot and two-shot all-reduce fused with the
1/world_size averaging, scatter+all-gather broadcast); there is no NCCL call on these paths.

Replaces ProcessGroupNCCL on the reference's hot path (/root/reference/dpp.py:21,39; SURVEY.md
§2.5(a) K2/K3/K4/K6, work items W2-W7).  The control plane -- exchanging the 64-byte IPC
handles -- goes through ``torch.distributed`` (``all_gather_object``), as the north star allows.

``EmulatedWorld`` runs the very same kernels with all "ranks" on one GPU in one cooperative
launch; it exists so the collective logic can be tested on a single-GPU box.
"""
from __future__ import annotations

import ctypes
import os
import struct
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from ..ops import native
from .comm import CommBase, _Done

_DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1}
HEADER_BYTES = 64 << 10          # signal pad + epoch counters + error word


class _RawMem:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr: int, nbytes: int) -> None:
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _as_tensor(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawMem(ptr, nbytes), device=device)


def _pack_ctx(bufs: Sequence[int], flags: Sequence[int], epochs: int, error: int, rank: int, world: int,
              channel: int, emulate: int, multicast: int = 0) -> ctypes.Array:
    b = list(bufs) + [0] * (8 - len(bufs))
    f = list(flags) + [0] * (8 - len(flags))
    raw = struct.pack("<8Q8QQQ4iQ", *b, *f, epochs, error, rank, world, channel, emulate, multicast)
    lib = native.load_library()
    assert len(raw) == lib.b200_comm_ctx_bytes(), (len(raw), lib.b200_comm_ctx_bytes())
    return ctypes.create_string_buffer(raw, len(raw))


def _lib():
    lib = native.load_library()
    for name in ("b200_comm_signal_bytes", "b200_comm_epoch_bytes", "b200_allreduce_oneshot_max_bytes"):
        getattr(lib, name).restype = ctypes.c_size_t
    return lib


class SymmetricArena:
    """One IPC-shared device block per rank with a bump allocator (same sequence on all ranks)."""

    def __init__(self, group, device: torch.device, nbytes: int, provider: str = "ipc") -> None:
        lib = _lib()
        self.group, self.device = group, device
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.nbytes = int(-(-nbytes // (2 << 20)) * (2 << 20))
        self.multicast = 0
        self.provider = provider
        assert lib.b200_comm_signal_bytes() + lib.b200_comm_epoch_bytes() + 256 <= HEADER_BYTES
        torch.cuda.set_device(device)
        if provider == "symm_mem":
            # PyTorch's symmetric-memory allocator (CUDA VMM + NVLink multicast object) as the memory
            # provider: same peer-pointer model, plus a multicast address for the NVLS kernels.  Only
            # allocation / handle exchange is delegated (control plane); all kernels are ours.
            import torch.distributed._symmetric_memory as symm_mem

            pg = group if group is not None else dist.group.WORLD
            symm_mem.enable_symm_mem_for_group(pg.group_name)
            self._symm_tensor = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
            hdl = symm_mem.rendezvous(self._symm_tensor, pg.group_name)
            self._symm_handle = hdl
            self.base = int(self._symm_tensor.data_ptr())
            self.peer_base = [int(p) for p in hdl.buffer_ptrs]
            self.multicast = int(hdl.multicast_ptr or 0)       # 0 when the box exposes no NVLink multicast
            self._opened = []
            self.memory = self._symm_tensor
            self.memory.zero_()
            self._finish_init(lib)
            return
        ptr = ctypes.c_void_p()
        rc = lib.b200_arena_alloc(ctypes.c_size_t(self.nbytes), ctypes.byref(ptr))
        if rc != 0:
            raise RuntimeError(f"arena cudaMalloc({self.nbytes}) failed: CUDA error {rc}")
        self.base = int(ptr.value)
        handle = (ctypes.c_ubyte * 64)()
        rc = lib.b200_ipc_get_handle(ctypes.c_void_p(self.base), handle)
        if rc != 0:
            raise RuntimeError(f"cudaIpcGetMemHandle failed: CUDA error {rc}")
        gathered: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(gathered, bytes(handle), group=group)
        self.peer_base: List[int] = []
        self._opened: List[int] = []
        for r, h in enumerate(gathered):
            if r == self.rank:
                self.peer_base.append(self.base)
                continue
            out = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            rc = lib.b200_ipc_open_handle(buf, ctypes.byref(out))
            if rc != 0:
                raise RuntimeError(f"cudaIpcOpenMemHandle(rank {r}) failed: CUDA error {rc}; "
                                   "peer access over NVLink is required for comm='p2p'")
            self.peer_base.append(int(out.value))
            self._opened.append(int(out.value))
        self.memory = _as_tensor(self.base, self.nbytes, device)      # uint8 view of the local arena
        self._finish_init(lib)

    def _finish_init(self, lib) -> None:
        group = self.group
        self._cursor = HEADER_BYTES
        sig = lib.b200_comm_signal_bytes()
        self.epochs_ptr = self.base + sig
        self.error_ptr = self.base + sig + lib.b200_comm_epoch_bytes()
        self._ctx: Dict[int, ctypes.Array] = {}
        dist.barrier(group=group)      # everyone has mapped everyone before the first kernel

    def ctx(self, channel: int) -> ctypes.Array:
        c = self._ctx.get(channel)
        if c is None:
            c = self._ctx[channel] = _pack_ctx(self.peer_base, self.peer_base, self.epochs_ptr, self.error_ptr,
                                               self.rank, self.world, channel, 0, self.multicast)
        return c

    def alloc(self, nbytes: int, align: int = 256) -> int:
        off = -(-self._cursor // align) * align
        end = off + -(-int(nbytes) // 16) * 16
        if end > self.nbytes:
            raise RuntimeError(f"symmetric arena exhausted: need {end} of {self.nbytes} bytes "
                               "(raise B200DDP_ARENA_MB)")
        self._cursor = end
        return off

    def view(self, offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        return self.memory[offset:offset + nbytes].view(dtype)

    def offset_of(self, t: torch.Tensor) -> Optional[int]:
        off = t.data_ptr() - self.base
        return off if 0 <= off and off + t.numel() * t.element_size() <= self.nbytes else None

    def error_word(self) -> int:
        off = self.error_ptr - self.base
        return int(self.memory[off:off + 4].view(torch.int32).item()) & 0xFFFFFFFF

    def close(self) -> None:
        lib = native.load_library()
        for p in self._opened:
            lib.b200_ipc_close_handle(ctypes.c_void_p(p))
        self._opened = []


class _EventHandle:
    def __init__(self, event: torch.cuda.Event) -> None:
        self.event = event

    def wait(self) -> None:
        torch.cuda.current_stream().wait_event(self.event)


class P2PComm(CommBase):
    """Gradient buckets reduced by our NVLink kernels on a high-priority side stream."""

    name = "p2p"
    GRAD_CHANNEL, MAIN_CHANNEL = 0, 1

    def __init__(self, process_group, device: torch.device, algo: str = "auto", arena_bytes: Optional[int] = None,
                 max_blocks: Optional[int] = None) -> None:
        super().__init__(process_group)
        self.device = device
        self.algo = algo
        self.max_blocks = int(max_blocks or os.environ.get("B200DDP_COMM_BLOCKS", 32))
        self.last_blocks = int(os.environ.get("B200DDP_COMM_BLOCKS_LAST", self.max_blocks))
        self.oneshot_max = int(os.environ.get("B200DDP_ONESHOT_MAX_BYTES", 256 << 10))
        self._arena_bytes = arena_bytes
        self.arena: Optional[SymmetricArena] = None
        self.stream = torch.cuda.Stream(device=device, priority=-1)
        self._staging: Optional[int] = None
        self._staging_bytes = 0
        self._buffer_region = None
        self.bytes_reduced = 0
        # observability / safety switches
        self.timing = os.environ.get("B200DDP_COMM_TIMING", "0") != "0"     # device-time every bucket reduction
        self.debug_check = os.environ.get("B200DDP_DEBUG_COMM", "0") != "0"  # cross-check each bucket against NCCL
        self.nvtx = os.environ.get("B200DDP_NVTX", "0") != "0"
        self._timed = []            # (start event, end event, bytes)
        self.debug_max_err = 0.0

    # -- arena ------------------------------------------------------------------------------
    def reserve(self, grad_bytes: int, buffer_bytes: int = 0) -> None:
        """Collective: size and create the arena (2 bucket generations + buffers + staging)."""
        if self.arena is not None:
            return
        env = os.environ.get("B200DDP_ARENA_MB")
        if self._arena_bytes is None:
            staging = min(max(grad_bytes, 16 << 20), 256 << 20)
            self._arena_bytes = (int(env) << 20) if env else int(2.3 * grad_bytes) + buffer_bytes + staging + (8 << 20)
        provider = os.environ.get("B200DDP_ARENA", "symm_mem" if self.algo == "nvls" else "ipc")
        self.arena = SymmetricArena(self.group, self.device, self._arena_bytes, provider=provider)
        if self.algo == "nvls" and not self.arena.multicast:
            import warnings

            warnings.warn("NVLS requested but this box exposes no NVLink multicast: using the two-shot P2P kernel")

    def _need_arena(self) -> SymmetricArena:
        if self.arena is None:
            self.reserve(256 << 20)
        return self.arena

    def allocate(self, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
        arena = self._need_arena()
        esize = torch.empty((), dtype=dtype).element_size()
        off = arena.alloc(int(numel) * esize)
        t = arena.view(off, int(numel), dtype)
        t.zero_()
        return t

    # -- collectives ------------------------------------------------------------------------
    def _pick_algo(self, nbytes: int) -> int:
        lib = _lib()
        if self.algo == "oneshot" and nbytes <= lib.b200_allreduce_oneshot_max_bytes():
            return 1
        if self.algo == "nvls":
            return 3 if (self.arena is not None and self.arena.multicast) else 2
        if self.algo == "twoshot":
            return 2
        if self.algo == "auto" and nbytes <= self.oneshot_max:
            return 1
        # with a multicast-capable arena (B200DDP_ARENA=symm_mem) the in-switch reduction wins from ~256 KB up
        # (8 GPUs, graph-timed: 44.7 MB 130 us vs 150 us two-shot; 256 MB 576 us vs 741 us)
        if self.algo == "auto" and self.arena is not None and self.arena.multicast and nbytes >= (256 << 10):
            return 3
        return 2

    def _allreduce_on(self, stream: torch.cuda.Stream, channel: int, off: int, nbytes: int, dtype: torch.dtype,
                      blocks: Optional[int] = None) -> None:
        native._call("b200_allreduce", self.arena.ctx(channel), ctypes.c_size_t(off), ctypes.c_size_t(nbytes),
                     ctypes.c_int(_DTYPE_CODE[dtype]), ctypes.c_float(1.0 / self.divisor),
                     ctypes.c_int(self._pick_algo(nbytes)), ctypes.c_int(blocks or self.max_blocks),
                     ctypes.c_void_p(stream.cuda_stream))

    def launch(self, index: int, flat: torch.Tensor, is_last: bool):
        self.launches += 1
        if self.world_size == 1:
            return _Done()
        off = self.arena.offset_of(flat) if self.arena is not None else None
        if off is None or flat.dtype not in _DTYPE_CODE:
            raise RuntimeError("P2PComm can only reduce buckets that live in the symmetric arena (fp32 / bf16)")
        nbytes = flat.numel() * flat.element_size()
        cur = torch.cuda.current_stream(self.device)
        capturing = torch.cuda.is_current_stream_capturing()
        if self.nvtx:
            torch.cuda.nvtx.range_push(f"b200ddp.bucket{index}.allreduce[{nbytes}B]")
        expect = None
        if self.debug_check and not capturing:
            expect = flat.detach().float().clone()
            dist.all_reduce(expect, group=self.group)
            expect /= self.divisor
        ready = torch.cuda.Event()
        ready.record(cur)
        self.stream.wait_event(ready)
        t0 = None
        if self.timing and not capturing:
            t0 = torch.cuda.Event(enable_timing=True)
            t0.record(self.stream)
        # the last bucket of a backward pass has nothing left to overlap with: it may use more of the chip
        # (every rank takes the same decision, so the block counts agree)
        self._allreduce_on(self.stream, self.GRAD_CHANNEL, off, nbytes, flat.dtype,
                           blocks=self.last_blocks if is_last else None)
        done = torch.cuda.Event(enable_timing=t0 is not None)
        done.record(self.stream)
        if t0 is not None:
            self._timed.append((t0, done, nbytes))
        self.bytes_reduced += nbytes
        if self.nvtx:
            torch.cuda.nvtx.range_pop()
        if expect is not None:
            done.synchronize()
            err = float((flat.float() - expect).abs().max() / expect.abs().max().clamp_min(1e-20))
            self.debug_max_err = max(self.debug_max_err, err)
            if err > (1e-2 if flat.dtype == torch.bfloat16 else 1e-4):
                raise RuntimeError(f"P2P all-reduce of bucket {index} disagrees with NCCL (max rel err {err:.3e})")
        return _EventHandle(done)

    def stats(self) -> dict:
        """Device-timed per-bucket statistics (needs B200DDP_COMM_TIMING=1; eager steps only)."""
        out = {"bucket_launches": self.launches, "bytes_reduced": self.bytes_reduced,
               "debug_max_rel_err_vs_nccl": self.debug_max_err}
        done = [(a.elapsed_time(b), n) for a, b, n in self._timed if b.query()]
        if done:
            ms = sum(t for t, _ in done)
            nbytes = sum(n for _, n in done)
            f = 2 * (self.world_size - 1) / self.world_size
            out.update(timed_buckets=len(done), avg_bucket_allreduce_us=1e3 * ms / len(done),
                       bus_bandwidth_GBs=nbytes / (ms * 1e-3) * f / 1e9 if ms > 0 else 0.0)
        return out

    def time_allreduce(self, nbytes: int, iters: int = 20, dtype: torch.dtype = torch.float32) -> dict:
        """Collective: device time of one in-place all-reduce of ``nbytes`` (kernel picked as for a gradient
        bucket), measured over ``iters`` back-to-back launches replayed from a CUDA graph, max over ranks.
        Bus bandwidth uses the 2(N-1)/N convention of the BASELINE metric."""
        arena = self._need_arena()
        nbytes = -(-int(nbytes) // 16) * 16
        self._ensure_staging()
        if nbytes > self._staging_bytes:
            return {"bytes": nbytes, "skipped": "larger than the staging region"}
        off = self._staging
        arena.memory[off:off + nbytes].zero_()
        algo = self._pick_algo(nbytes)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._allreduce_on(side, self.MAIN_CHANNEL, off, nbytes, dtype)
            side.synchronize()
            dist.barrier(group=self.group)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
                for _ in range(iters):
                    self._allreduce_on(side, self.MAIN_CHANNEL, off, nbytes, dtype)
            graph.replay()
            side.synchronize()
            dist.barrier(group=self.group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(side)
            graph.replay()
            e1.record(side)
            side.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        us = float(t.item()) * 1e3
        f = 2 * (self.world_size - 1) / self.world_size
        return {"bytes": nbytes, "us": round(us, 2), "bus_gbs": round(nbytes * f / (us * 1e-6) / 1e9, 1),
                "algo": {1: "oneshot", 2: "twoshot", 3: "nvls"}[algo]}

    def all_reduce_avg_(self, flat: torch.Tensor) -> None:
        """In-place average on the current stream (tensor inside the arena, or staged through it)."""
        cur = torch.cuda.current_stream(self.device)
        off = self.arena.offset_of(flat) if self.arena is not None else None
        if off is not None and flat.is_contiguous() and (flat.numel() * flat.element_size()) % 16 == 0:
            self._allreduce_on(cur, self.MAIN_CHANNEL, off, flat.numel() * flat.element_size(), flat.dtype)
            return
        for view, stage_off, nbytes in self._staged_chunks(flat):
            stage = self.arena.memory[stage_off:stage_off + view.numel() * view.element_size()].view(flat.dtype)
            pad = nbytes - view.numel() * view.element_size()
            stage.copy_(view)
            if pad:
                self.arena.memory[stage_off + nbytes - pad:stage_off + nbytes].zero_()
            self._allreduce_on(cur, self.MAIN_CHANNEL, stage_off, nbytes, flat.dtype)
            view.copy_(stage)

    def _ensure_staging(self) -> None:
        arena = self._need_arena()
        if self._staging is None:
            self._staging_bytes = min(64 << 20, max(1 << 20, (arena.nbytes - arena._cursor) // 2))
            self._staging = arena.alloc(self._staging_bytes)

    def _staged_chunks(self, flat: torch.Tensor):
        self._ensure_staging()
        flat1 = flat.reshape(-1)
        esize = flat.element_size()
        per = self._staging_bytes // esize
        for s in range(0, flat1.numel(), per):
            view = flat1[s:s + per]
            nbytes = -(-view.numel() * esize // 16) * 16
            yield view, self._staging, nbytes

    def broadcast_(self, flat: torch.Tensor, src: int = 0) -> None:
        if self.world_size == 1:
            return
        cur = torch.cuda.current_stream(self.device)
        off = self.arena.offset_of(flat) if self.arena is not None else None
        nb = flat.numel() * flat.element_size()
        if off is not None and flat.is_contiguous() and nb % 16 == 0 and off % 16 == 0:
            self._broadcast_on(cur, off, nb, src)
            return
        if not flat.is_contiguous():
            raise ValueError("broadcast_ needs a contiguous tensor")
        for view, stage_off, nbytes in self._staged_chunks(flat):
            stage = self.arena.memory[stage_off:stage_off + view.numel() * view.element_size()].view(flat.dtype)
            if self.rank == src:
                stage.copy_(view)
            self._broadcast_on(cur, stage_off, nbytes, src)
            if self.rank != src:
                view.copy_(stage)

    def _broadcast_on(self, stream, off: int, nbytes: int, src: int) -> None:
        native._call("b200_broadcast", self.arena.ctx(self.MAIN_CHANNEL), ctypes.c_size_t(off),
                     ctypes.c_size_t(nbytes), ctypes.c_int(src), ctypes.c_int(self.max_blocks),
                     ctypes.c_void_p(stream.cuda_stream))

    def barrier(self) -> None:
        if self.world_size > 1:
            native._call("b200_barrier", self._need_arena().ctx(self.MAIN_CHANNEL),
                         ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))

    # -- per-forward module-buffer sync (K3): one launch for every buffer of every dtype --------
    def sync_buffers(self, module: torch.nn.Module, buffers: Sequence[torch.Tensor]) -> int:
        if self.world_size == 1 or not buffers:
            return 0
        if self._buffer_region is None:
            arena = self._need_arena()
            total = sum(-(-b.numel() * b.element_size() // 16) * 16 for b in buffers)
            base = arena.alloc(total)
            off = base
            with torch.no_grad():
                for b in buffers:
                    view = arena.view(off, b.numel(), b.dtype).view(b.shape)
                    view.copy_(b)
                    b.data = view                      # the module's buffer now lives in the arena
                    off += -(-b.numel() * b.element_size() // 16) * 16
            self._buffer_region = (base, total)
        base, total = self._buffer_region
        self._broadcast_on(torch.cuda.current_stream(self.device), base, total, 0)
        return 1

    def check(self) -> None:
        """Raise if a kernel gave up waiting for a peer (bounded spin)."""
        if self.arena is not None:
            torch.cuda.synchronize(self.device)
            w = self.arena.error_word()
            if w:
                raise RuntimeError(f"P2P collective timed out waiting for rank {w & 0xFFFF} (error word {w:#x})")


class EmulatedWorld:
    """All ranks on ONE GPU: same kernels, one cooperative launch with blockIdx.y as the rank."""

    def __init__(self, world: int, device: torch.device | str = "cuda", arena_bytes: int = 64 << 20,
                 max_blocks: int = 8) -> None:
        lib = _lib()
        self.world, self.device = world, torch.device(device)
        self.nbytes = arena_bytes
        self.max_blocks = max_blocks
        self.arenas = [torch.zeros(arena_bytes, dtype=torch.uint8, device=self.device) for _ in range(world)]
        self.epochs = torch.zeros(lib.b200_comm_epoch_bytes() // 4, dtype=torch.int32, device=self.device)
        self.error = torch.zeros(64, dtype=torch.int32, device=self.device)
        self.pads = [torch.zeros(lib.b200_comm_signal_bytes() // 4, dtype=torch.int32, device=self.device)
                     for _ in range(world)]
        bases = [a.data_ptr() for a in self.arenas]
        pads = [p.data_ptr() for p in self.pads]
        self._ctx = {ch: _pack_ctx(bases, pads, self.epochs.data_ptr(), self.error.data_ptr(), 0, world, ch, 1)
                     for ch in range(2)}

    def view(self, rank: int, offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        return self.arenas[rank][offset:offset + nbytes].view(dtype)

    def allreduce(self, offset: int, nbytes: int, dtype: torch.dtype, algo: int, scale: Optional[float] = None,
                  channel: int = 0) -> None:
        native._call("b200_allreduce", self._ctx[channel], ctypes.c_size_t(offset), ctypes.c_size_t(nbytes),
                     ctypes.c_int(_DTYPE_CODE[dtype]), ctypes.c_float(1.0 / self.world if scale is None else scale),
                     ctypes.c_int(algo), ctypes.c_int(self.max_blocks),
                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def broadcast(self, offset: int, nbytes: int, root: int, channel: int = 1) -> None:
        native._call("b200_broadcast", self._ctx[channel], ctypes.c_size_t(offset), ctypes.c_size_t(nbytes),
                     ctypes.c_int(root), ctypes.c_int(self.max_blocks),
                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def barrier(self, channel: int = 1) -> None:
        native._call("b200_barrier", self._ctx[channel], ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def error_word(self) -> int:
        return int(self.error[0].item()) & 0xFFFFFFFF


class FusedGradReduce:
    """Weight-gradient GEMM whose epilogue is the all-reduce.

    ``dW = average_over_ranks(dY_r^T @ X_r)`` computed by ONE tcgen05 kernel per rank: partial
    tiles are pushed over NVLink to their owner rank, the owner sums them in rank order and pushes
    the finished tile into every rank's copy of ``out`` (which must live in the symmetric arena,
    e.g. a gradient-bucket slot), so the transfer of one tile overlaps the MMA of the next.
    """

    BM = 128

    def __init__(self, comm: "P2PComm") -> None:
        self.comm = comm
        self._stage = None          # (offset, bytes)
        self._flags = None          # (offset, words)
        self._epoch = None
        lib = _lib()
        assert lib.b200_peer_reduce_bytes() == 216, lib.b200_peer_reduce_bytes()

    def _reserve(self, tiles: int, bn: int) -> None:
        arena = self.comm.arena
        world = self.comm.world_size
        need = world * tiles * self.BM * bn * 4
        if self._stage is None or self._stage[1] < need:
            self._stage = (arena.alloc(need), need)
        words = tiles * 9
        if self._flags is None or self._flags[1] < words:
            off = arena.alloc(words * 4)
            arena.memory[off:off + words * 4].zero_()
            self._flags = (off, words)
        if self._epoch is None:
            self._epoch = arena.alloc(256)
            arena.memory[self._epoch:self._epoch + 256].zero_()

    def wgrad(self, dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor) -> None:
        """out[N_out, K_in] (arena resident, same offset on every rank) = mean_r(dy_r^T @ x_r)."""
        arena = self.comm.arena
        off = arena.offset_of(out)
        if off is None or not out.is_contiguous():
            raise RuntimeError("fused gradient reduction needs a contiguous output inside the symmetric arena")
        m_tok, n_out = dy.shape
        k_in = x.shape[1]
        lib = _lib()
        bn = lib.b200_gemm_tc_peer_bn(ctypes.c_int(k_in))
        tiles = -(-n_out // self.BM) * -(-k_in // bn)
        self._reserve(tiles, bn)
        world, rank = self.comm.world_size, self.comm.rank
        pb = arena.peer_base
        pad = [0] * (8 - world)
        raw = struct.pack("<8Q8Q8QQiifi",
                          *[b + off for b in pb], *pad,
                          *[b + self._stage[0] for b in pb], *pad,
                          *[b + self._flags[0] for b in pb], *pad,
                          arena.base + self._epoch, rank, world, 1.0 / world, 0)
        peer = ctypes.create_string_buffer(raw, len(raw))
        flags = int(dy.dtype == torch.bfloat16) | (int(out.dtype == torch.bfloat16) << 1)
        desc = (ctypes.c_longlong * 11)(n_out, k_in, k_in, 1, 1, 0, 0, 0, 0, 0, 0)
        err = native.scratch(dy.device).error
        native._call("b200_gemm_tc_peer", ctypes.c_void_p(dy.data_ptr()), ctypes.c_void_p(x.data_ptr()),
                     ctypes.c_void_p(0), ctypes.c_int(n_out), ctypes.c_int(k_in), ctypes.c_int(m_tok), ctypes.c_int(2),
                     ctypes.c_int(flags), desc, peer, ctypes.c_void_p(err.data_ptr()),
                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), kernels=2)
