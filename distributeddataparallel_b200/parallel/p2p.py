"""Symmetric arena + the P2P (NVLink) communication backend.

This is the product data plane: gradient buckets, module buffers and broadcast / push staging live in
a per-rank *symmetric arena* -- one device block per rank that every peer has mapped (the C++ ``_C.VmmArena``:
``cuMemCreate`` + fd exchange + NVLink multicast bind; CUDA IPC as the fallback), so that an object sits at the same
offset in every rank's arena and each rank can ``ld``/``st`` its peers' copies directly over NVLink 5 / NVSwitch,
or reduce / broadcast them inside the switch through the multicast mapping.  The collectives are the kernels in
``csrc/kernels/comm.cu`` (barrier; push, one-shot, two-shot and NVLS all-reduce fused with the 1/world_size
averaging; a 16-bit-wire form for the compress hooks; scatter+all-gather broadcast); there is no NCCL call on
these paths.

Replaces ProcessGroupNCCL on the reference's hot path (/root/reference/dpp.py:21,39; SURVEY.md
§2.5(a) K2/K3/K4/K6, work items W2-W7).  The control plane -- agreeing on the outcome of every set-up stage,
naming the socket namespace -- goes through ``torch.distributed`` (``all_gather_object``), as the north star allows.

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
from .comm import CommBase, _Done, _EventHandle

_DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1}
HEADER_BYTES = 64 << 10          # signal pad + epoch counters + error word


class _RawMem:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr: int, nbytes: int) -> None:
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _as_tensor(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawMem(ptr, nbytes), device=device)


def _pack_ctx(bufs: Sequence[int], flags: Sequence[int], epochs: int, error: int, rank: int, world: int,
              channel: int, emulate: int, multicast: int = 0, timeout_ns: int = 0, stamps: int = 0) -> ctypes.Array:
    b = list(bufs) + [0] * (8 - len(bufs))
    f = list(flags) + [0] * (8 - len(flags))
    raw = struct.pack("<8Q8QQQ4iQQQ", *b, *f, epochs, error, rank, world, channel, emulate, multicast, timeout_ns, stamps)
    lib = native.load_library()
    assert len(raw) == lib.b200_comm_ctx_bytes(), (len(raw), lib.b200_comm_ctx_bytes())
    return ctypes.create_string_buffer(raw, len(raw))


def _lib():
    lib = native.load_library()
    for name in ("b200_comm_signal_bytes", "b200_comm_epoch_bytes", "b200_allreduce_oneshot_max_bytes",
                 "b200_allreduce_push_max_bytes", "b200_allreduce_push_stage_bytes"):
        getattr(lib, name).restype = ctypes.c_size_t
    return lib


def comm_timeout_ns() -> int:
    """Bound on every in-kernel wait for a peer (``B200DDP_COMM_TIMEOUT_S``, default 600 s: the NCCL
    process-group timeout the reference inherits through /root/reference/dpp.py:21)."""
    return int(float(os.environ.get("B200DDP_COMM_TIMEOUT_S", "600")) * 1e9)


class ArenaUnavailable(RuntimeError):
    """Raised by EVERY rank together when the symmetric arena cannot be set up on some rank."""


def _agree(group, local_error: str) -> Optional[str]:
    """Collective: every rank contributes "" or an error; returns None when all succeeded, else one message
    (the same on every rank), so that all ranks take the same branch."""
    world = dist.get_world_size(group)
    gathered: List[Optional[str]] = [None] * world
    dist.all_gather_object(gathered, local_error or "", group=group)
    bad = [(r, e) for r, e in enumerate(gathered) if e]
    return None if not bad else f"rank {bad[0][0]}: {bad[0][1]}"


class SymmetricArena:
    """One device block per rank, mapped into every peer, with a bump allocator (same sequence on all ranks).

    Providers (``B200DDP_ARENA``):

    * ``vmm`` (default): this framework's own allocator in C++ (``csrc/runtime/vmm_arena.cpp``) -- ``cuMemCreate`` +
      POSIX-fd export over a unix socket + ``cuMemMap`` of every peer, plus an NVLink *multicast* object all ranks
      bind to (``cuMulticastCreate/AddDevice/BindMem``) where the device supports it, which enables the NVLS kernels;
    * ``ipc``: ``cudaMalloc`` + CUDA IPC handles (no multicast); the fallback when the VMM path is unavailable.

    Set-up is staged; after every stage the ranks agree on the outcome, so a failure on one rank makes ALL ranks
    raise :class:`ArenaUnavailable` at the same point instead of leaving peers inside a collective.
    """

    def __init__(self, group, device: torch.device, nbytes: int, provider: str = "vmm") -> None:
        lib = _lib()
        self.group, self.device = group, device
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.nbytes = int(-(-nbytes // (2 << 20)) * (2 << 20))
        self.multicast = 0
        self.provider = provider
        self._vmm = None
        self._opened: List[int] = []
        self._ipc_base = 0
        assert lib.b200_comm_signal_bytes() + lib.b200_comm_epoch_bytes() + 256 <= HEADER_BYTES
        torch.cuda.set_device(device)
        torch.zeros(1, device=device)                 # make sure the primary context exists and is current
        if provider == "vmm":
            err = self._init_vmm()
            if err is not None:
                if os.environ.get("B200DDP_ARENA") == "vmm":
                    raise ArenaUnavailable(f"VMM arena unavailable ({err})")
                self.provider = provider = "ipc"
                self.vmm_error = err
        if provider == "ipc":
            err = self._init_ipc(lib)
            if err is not None:
                raise ArenaUnavailable(f"CUDA IPC arena unavailable ({err})")
        elif provider != "vmm":
            raise ValueError(f"unknown arena provider {provider!r}")
        self.memory = _as_tensor(self.base, self.nbytes, device)      # uint8 view of the local arena
        self.memory.zero_()
        torch.cuda.synchronize(device)
        self._finish_init(lib)

    # -- providers ---------------------------------------------------------------------------
    def _init_vmm(self) -> Optional[str]:
        from .. import _C

        tokens: List[Optional[str]] = [None] * self.world
        dist.all_gather_object(tokens, os.urandom(6).hex(), group=self.group)
        session = f"{os.environ.get('MASTER_PORT', '0')}-{tokens[0]}"       # rank 0's token names the socket namespace
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        vmm = _C.VmmArena(idx, self.nbytes, self.rank, self.world, session)
        stages = (("alloc", vmm.alloc_local), ("fd exchange", vmm.exchange), ("map peers", vmm.map_peers))
        for name, fn in stages:
            try:
                local = fn()
            except Exception as e:  # pragma: no cover
                local = repr(e)
            err = _agree(self.group, f"{name}: {local}" if local else "")
            if err is not None:
                vmm.close()
                return err
        self._vmm = vmm
        self.nbytes = int(vmm.nbytes())
        self.base = int(vmm.base())
        self.peer_base = [int(vmm.peer_base(r)) for r in range(self.world)]
        # NVLink multicast (NVLS): optional -- every rank must succeed at every stage or nobody uses it
        if os.environ.get("B200DDP_MULTICAST", "1") != "0":
            err = _agree(self.group, vmm.multicast_create())
            if err is None:
                err = _agree(self.group, vmm.multicast_bind())
                if err is None:
                    self.multicast = int(vmm.multicast_base())
            self.multicast_error = err
        return None

    def _init_ipc(self, lib) -> Optional[str]:
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        local = ""
        rc = lib.b200_arena_alloc(ctypes.c_size_t(self.nbytes), ctypes.byref(ptr))
        if rc != 0:
            local = f"cudaMalloc({self.nbytes}) failed: CUDA error {rc}"
        else:
            self._ipc_base = int(ptr.value)
            rc = lib.b200_ipc_get_handle(ctypes.c_void_p(self._ipc_base), handle)
            if rc != 0:
                local = f"cudaIpcGetMemHandle failed: CUDA error {rc}"
        gathered: List[Optional[tuple]] = [None] * self.world
        dist.all_gather_object(gathered, (local, bytes(handle)), group=self.group)
        bad = [(r, e) for r, (e, _) in enumerate(gathered) if e]
        if bad:
            self._free_ipc(lib)
            return f"rank {bad[0][0]}: {bad[0][1]}"
        self.base = self._ipc_base
        self.peer_base = []
        local = ""
        for r, (_, h) in enumerate(gathered):
            if r == self.rank:
                self.peer_base.append(self.base)
                continue
            out = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            rc = lib.b200_ipc_open_handle(buf, ctypes.byref(out))
            if rc != 0:
                local = local or (f"cudaIpcOpenMemHandle(rank {r}) failed: CUDA error {rc}; peer access over "
                                  "NVLink is required for comm='p2p'")
                self.peer_base.append(0)
                continue
            self.peer_base.append(int(out.value))
            self._opened.append(int(out.value))
        err = _agree(self.group, local)
        if err is not None:
            self._free_ipc(lib)
        return err

    def _free_ipc(self, lib) -> None:
        for p in self._opened:
            lib.b200_ipc_close_handle(ctypes.c_void_p(p))
        self._opened = []
        if self._ipc_base:
            lib.b200_arena_free(ctypes.c_void_p(self._ipc_base))
            self._ipc_base = 0

    def _finish_init(self, lib) -> None:
        group = self.group
        self._cursor = HEADER_BYTES
        sig = lib.b200_comm_signal_bytes()
        self.epochs_ptr = self.base + sig
        # the error word lives in pinned host memory: kernels store to it over PCIe only when a wait times out,
        # the host reads it with a plain load -- cheap enough to poll after every step
        self._error_host = torch.zeros(16, dtype=torch.int32).pin_memory()
        self.error_ptr = self._error_host.data_ptr()
        self.timeout_ns = comm_timeout_ns()
        self._ctx: Dict[int, ctypes.Array] = {}
        dist.barrier(group=group)      # everyone has mapped everyone before the first kernel

    def ctx(self, channel: int, stamps: int = 0) -> ctypes.Array:
        """Host copy of the kernels' context for one channel; ``stamps`` = device address of a (start, end) pair of
        64-bit words the collective writes %globaltimer into (device-timed logger), 0 for none."""
        c = self._ctx.get((channel, stamps))
        if c is None:
            c = self._ctx[(channel, stamps)] = _pack_ctx(self.peer_base, self.peer_base, self.epochs_ptr, self.error_ptr,
                                                         self.rank, self.world, channel, 0, self.multicast,
                                                         self.timeout_ns, stamps)
        return c

    def set_timeout(self, seconds: float) -> None:
        self.timeout_ns = int(seconds * 1e9)
        self._ctx.clear()

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
        """Host-side read of the error word (no CUDA call, no synchronisation)."""
        return int(self._error_host[0]) & 0xFFFFFFFF

    def seed_epochs(self, value: int) -> None:
        """Test hook: start every epoch counter (and every flag, on every rank -- call it on all ranks, then
        barrier) at ``value``, e.g. 2**32 - 8 to exercise the wraparound."""
        lib = _lib()
        sig, ep = lib.b200_comm_signal_bytes(), lib.b200_comm_epoch_bytes()
        v = value & 0xFFFFFFFF
        v = v - (1 << 32) if v >= (1 << 31) else v
        self.memory[:sig + ep].view(torch.int32).fill_(v)
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        lib = native.load_library()
        if self._vmm is not None:
            torch.cuda.synchronize(self.device)
            self._vmm.close()
            self._vmm = None
        else:
            for p in self._opened:
                lib.b200_ipc_close_handle(ctypes.c_void_p(p))
            self._opened = []


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
        # Gradient buckets reduced UNDER the backward pass use few CTAs: the reductions are switch / link bound, not
        # SM bound, and every CTA they hold slows the latency-bound backward kernels they share the chip with
        # (8 GPUs, NVLS, dpp.py step: 32 CTAs 0.820 ms, 16 CTAs 0.797 ms, 8 CTAs 0.785 ms per step and the reductions
        # themselves did not get slower; the SM-driven two-shot kernel needs 16: 2 GPUs 0.768 ms vs 0.775 with 32 and
        # 0.830 with 8).  Stand-alone collectives (all_reduce_avg_, the sweep) keep `max_blocks`.
        self.overlap_blocks = int(os.environ.get("B200DDP_COMM_BLOCKS_OVERLAP", 16))
        self.nvls_overlap_blocks = int(os.environ.get("B200DDP_NVLS_BLOCKS_OVERLAP", 8))
        self.nvls_blocks = int(os.environ.get("B200DDP_NVLS_BLOCKS", 24))     # the switch reduces: few SMs suffice
        # algorithm boundaries, from the graph-timed sweeps on B200 (profiles/allreduce_sweep_n{2,4,8}.json):
        #   push (one flag exchange)  wins up to 256 KB on 2 ranks, 16 KB on 4, 8 KB on 8 (its (N-1) stores per vector grow
        #                             with the rank count);
        #   NVLS (in-switch)          wins from 256 KB on 8 ranks, only from 128 MB on 4, never on 2;
        #   two-shot                  everything in between.
        ws = self.world_size
        self.push_auto_max = int(os.environ.get("B200DDP_PUSH_MAX_BYTES",
                                                (256 << 10) if ws <= 2 else (16 << 10) if ws <= 4 else (8 << 10)))
        self.nvls_min_world = int(os.environ.get("B200DDP_NVLS_MIN_WORLD", 3))
        self.nvls_min_bytes = int(os.environ.get("B200DDP_NVLS_MIN_BYTES", (128 << 20) if ws <= 5 else (256 << 10)))
        self.oneshot_max = int(os.environ.get("B200DDP_ONESHOT_MAX_BYTES", 256 << 10))
        self._arena_bytes = arena_bytes
        self.arena: Optional[SymmetricArena] = None
        self.stream = torch.cuda.Stream(device=device, priority=int(os.environ.get("B200DDP_COMM_PRIORITY", "-1")))
        self._staging: Optional[int] = None
        self._staging_bytes = 0
        self._push_stage: Dict[int, int] = {}      # channel -> arena offset of the push all-reduce staging object
        self._buffer_region = None
        self._ahead_event: Optional[torch.cuda.Event] = None
        # device-timed logger: (start, end) %globaltimer pair per launch unit + one for the buffer broadcast
        self.MAX_STAMPED = 64
        self._stamps = torch.zeros(2 * (self.MAX_STAMPED + 1), dtype=torch.int64, device=device)
        self._stamp_bytes: Dict[int, int] = {}
        self.bytes_reduced = 0
        self.algo_counts: Dict[str, int] = {}
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
            self._arena_bytes = (int(env) << 20) if env else int(2.3 * grad_bytes) + buffer_bytes + staging + (20 << 20)
        provider = os.environ.get("B200DDP_ARENA", "vmm")
        self.arena = SymmetricArena(self.group, self.device, self._arena_bytes, provider=provider)
        lib = _lib()
        self.push_max = int(lib.b200_allreduce_push_max_bytes())
        for ch in (self.GRAD_CHANNEL, self.MAIN_CHANNEL):
            self._push_stage[ch] = self.arena.alloc(int(lib.b200_allreduce_push_stage_bytes()))
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
    ALGO_NAMES = {1: "oneshot", 2: "twoshot", 3: "nvls", 4: "push"}

    def _pick_algo(self, nbytes: int) -> int:
        """1 one-shot, 2 two-shot, 3 NVLS (multimem), 4 push (single flag exchange).  ``auto``: push up to 256 KB
        (latency: one exchange instead of two barriers), NVLS above that wherever the arena is multicast-bound (the
        switch reduces: measured faster than two-shot from 1 MB up on 8 GPUs), else two-shot."""
        lib = _lib()
        if self.algo == "oneshot" and nbytes <= lib.b200_allreduce_oneshot_max_bytes():
            return 1
        if self.algo == "push" and nbytes <= self.push_max:
            return 4
        if self.algo == "nvls":
            return 3 if (self.arena is not None and self.arena.multicast) else 2
        if self.algo == "twoshot":
            return 2
        if self.algo == "auto":
            if nbytes <= min(self.push_auto_max, self.push_max):
                return 4
            # NVLS pays off where the switch saves link traffic: from 4 ranks up (on 2 ranks the two-shot kernel moves
            # S/2 per direction and is faster: 32 MB 67 us vs 106 us measured)
            if self.arena is not None and self.arena.multicast and self.world_size >= self.nvls_min_world \
                    and nbytes >= self.nvls_min_bytes:
                return 3
        return 2

    def _stamp_ptr(self, unit: Optional[int]) -> int:
        if unit is None or not (0 <= unit <= self.MAX_STAMPED):
            return 0
        return self._stamps.data_ptr() + 16 * unit

    def _allreduce_on(self, stream: torch.cuda.Stream, channel: int, off: int, nbytes: int, dtype: torch.dtype,
                      blocks: Optional[int] = None, algo: Optional[int] = None, unit: Optional[int] = None) -> int:
        algo = self._pick_algo(nbytes) if algo is None else algo
        stamps = self._stamp_ptr(unit)
        if algo == 4:
            native._call("b200_allreduce_push", self.arena.ctx(channel, stamps), ctypes.c_size_t(off), ctypes.c_size_t(nbytes),
                         ctypes.c_int(_DTYPE_CODE[dtype]), ctypes.c_float(1.0 / self.divisor),
                         ctypes.c_size_t(self._push_stage[channel]), ctypes.c_void_p(stream.cuda_stream))
            return algo
        if algo == 3:
            blocks = min(blocks or self.nvls_blocks, self.nvls_blocks)
        native._call("b200_allreduce", self.arena.ctx(channel, stamps), ctypes.c_size_t(off), ctypes.c_size_t(nbytes),
                     ctypes.c_int(_DTYPE_CODE[dtype]), ctypes.c_float(1.0 / self.divisor),
                     ctypes.c_int(algo), ctypes.c_int(blocks or self.max_blocks),
                     ctypes.c_void_p(stream.cuda_stream))
        return algo

    def launch(self, index: int, flat: torch.Tensor, is_last: bool, wire: Optional[torch.dtype] = None):
        """Reduce one launch unit on the communication stream.  ``wire`` = bf16 / fp16: fp32 data travels in that
        16-bit format and is accumulated in fp32 (the fused form of the compress comm hooks)."""
        self.launches += 1
        if self.world_size == 1:
            native.join_side_streams(self.device)
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
        if self.debug_check and not capturing and wire is None:
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
        for ev in native.side_stream_events():        # weight gradients produced on the side streams (ops/native.py)
            self.stream.wait_event(ev)
        if wire is not None and flat.dtype == torch.float32 and nbytes % 32 == 0 and nbytes > self.push_max:
            self.allreduce_wire16_(flat, wire, stream=self.stream, channel=self.GRAD_CHANNEL, unit=index)
            name = "wire16"
        else:
            algo = self._pick_algo(nbytes)
            blocks = self.last_blocks if is_last else (self.nvls_overlap_blocks if algo == 3 else self.overlap_blocks)
            self._allreduce_on(self.stream, self.GRAD_CHANNEL, off, nbytes, flat.dtype, blocks=blocks, algo=algo,
                               unit=index)
            name = self.ALGO_NAMES[algo]
        if 0 <= index < self.MAX_STAMPED:
            self._stamp_bytes[index] = nbytes
        self.algo_counts[name] = self.algo_counts.get(name, 0) + 1
        if self.after_reduce is not None and expect is None:
            with torch.cuda.stream(self.stream):
                self.after_reduce(index)            # e.g. the optimizer update of this unit, behind its reduction
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
               "debug_max_rel_err_vs_nccl": self.debug_max_err, "algo_counts": dict(self.algo_counts),
               "arena_provider": self.arena.provider if self.arena is not None else None,
               "multicast": bool(self.arena is not None and self.arena.multicast)}
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
                "algo": self.ALGO_NAMES[algo]}

    def raw_stamps(self):
        """(stamps, bytes, broadcast pair): the (start, end) %globaltimer pairs of every launch unit as a flat list, the
        payload bytes per unit, and the pair of the buffer broadcast.  Synchronises."""
        torch.cuda.synchronize(self.device)
        st = self._stamps.cpu().tolist()
        n = (max(self._stamp_bytes) + 1) if self._stamp_bytes else 0
        nbytes = [self._stamp_bytes.get(i, 0) for i in range(n)]
        return st[:2 * n], nbytes, (st[2 * self.MAX_STAMPED], st[2 * self.MAX_STAMPED + 1])

    def device_times(self) -> dict:
        """Device-side (start, end) of the last reduction of every launch unit and of the buffer broadcast, in ns of
        %globaltimer (written by the kernels themselves: valid under CUDA-graph replay).  Synchronises."""
        torch.cuda.synchronize(self.device)
        st = self._stamps.cpu().tolist()
        f = 2 * (self.world_size - 1) / max(self.world_size, 1)
        units = []
        for i in sorted(self._stamp_bytes):
            a, b = st[2 * i], st[2 * i + 1]
            if a and b and b >= a:
                nb = self._stamp_bytes[i]
                units.append({"unit": i, "bytes": nb, "start_ns": a, "end_ns": b, "us": (b - a) / 1e3,
                              "bus_gbs": nb * f / max(b - a, 1)})
        out = {"units": units}
        a, b = st[2 * self.MAX_STAMPED], st[2 * self.MAX_STAMPED + 1]
        if a and b and b >= a:
            out["buffer_broadcast"] = {"start_ns": a, "end_ns": b, "us": (b - a) / 1e3}
        return out

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
    def _buffers_into_arena(self, buffers: Sequence[torch.Tensor]) -> None:
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

    def sync_buffers(self, module: torch.nn.Module, buffers: Sequence[torch.Tensor]) -> int:
        if self.world_size == 1 or not buffers:
            return 0
        self._buffers_into_arena(buffers)
        base, total = self._buffer_region
        self._broadcast_on(torch.cuda.current_stream(self.device), base, total, 0)
        return 1

    def sync_buffers_ahead(self, module: torch.nn.Module, buffers: Sequence[torch.Tensor]) -> int:
        """K3 for the NEXT forward, issued on the communication stream right after this forward: rank 0's
        buffers are final once its forward has run, so the broadcast rides under the backward pass instead of
        heading the next step.  ``wait_buffers_ahead`` orders the next forward after it."""
        if self.world_size == 1 or not buffers:
            return 0
        self._buffers_into_arena(buffers)
        base, total = self._buffer_region
        cur = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(cur)
        self.stream.wait_event(ready)
        native._call("b200_broadcast", self.arena.ctx(self.GRAD_CHANNEL, self._stamp_ptr(self.MAX_STAMPED)),
                     ctypes.c_size_t(base), ctypes.c_size_t(total), ctypes.c_int(0), ctypes.c_int(self.max_blocks),
                     ctypes.c_void_p(self.stream.cuda_stream))
        self._ahead_event = torch.cuda.Event()
        self._ahead_event.record(self.stream)
        return 1

    def wait_buffers_ahead(self) -> None:
        ev, self._ahead_event = self._ahead_event, None
        if ev is not None and not torch.cuda.is_current_stream_capturing():
            torch.cuda.current_stream(self.device).wait_event(ev)

    def join_stream(self) -> None:
        """Order the current stream after everything issued on the communication stream (used before a capture
        ends and by the optimizer when no gradient bucket handle carries the dependency)."""
        ev = torch.cuda.Event()
        ev.record(self.stream)
        torch.cuda.current_stream(self.device).wait_event(ev)

    # -- compressed wire format (comm hooks) -----------------------------------------------------------------
    def allreduce_wire16_(self, flat: torch.Tensor, wire: torch.dtype, stream: Optional[torch.cuda.Stream] = None,
                          channel: Optional[int] = None, unit: Optional[int] = None) -> None:
        """In-place average of an fp32 arena tensor with bf16 / fp16 on the wire and fp32 accumulation."""
        off = self.arena.offset_of(flat)
        nbytes = flat.numel() * flat.element_size()
        if off is None or flat.dtype != torch.float32 or nbytes % 32 != 0:
            raise RuntimeError("wire16 all-reduce needs an fp32 arena tensor whose size is a multiple of 32 bytes")
        self._ensure_staging()
        if nbytes // 2 > self._staging_bytes:
            raise RuntimeError("bucket too large for the staging region")
        stream = stream or torch.cuda.current_stream(self.device)
        native._call("b200_allreduce_wire16", self.arena.ctx(self.MAIN_CHANNEL if channel is None else channel,
                                                             self._stamp_ptr(unit)),
                     ctypes.c_size_t(off), ctypes.c_size_t(nbytes), ctypes.c_int(1 if wire == torch.bfloat16 else 2),
                     ctypes.c_float(1.0 / self.divisor), ctypes.c_size_t(self._staging),
                     ctypes.c_int(self.max_blocks), ctypes.c_void_p(stream.cuda_stream))

    # -- failure detection -------------------------------------------------------------------------------------
    def check(self, sync: bool = True) -> None:
        """Raise if a kernel gave up waiting for a peer (time-bounded waits).  The error word lives
        in pinned host memory, so ``sync=False`` costs one host load and can run after every step."""
        if self.arena is not None:
            if sync:
                torch.cuda.synchronize(self.device)
            w = self.arena.error_word()
            if w:
                raise RuntimeError(
                    f"P2P collective timed out on rank {self.rank} waiting for rank {w & 0xFFFF} (error word {w:#x}, "
                    f"limit {self.arena.timeout_ns / 1e9:.0f} s): the affected results were poisoned with NaN; a peer is "
                    "dead or stalled -- aborting this rank")

    def close(self) -> None:
        if self.arena is not None:
            self.arena.close()
            self.arena = None


class EmulatedWorld:
    """All ranks on ONE GPU: same kernels, one cooperative launch with blockIdx.y as the rank."""

    def __init__(self, world: int, device: torch.device | str = "cuda", arena_bytes: int = 64 << 20,
                 max_blocks: int = 8, timeout_s: float = 30.0) -> None:
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
        self._ctx = {ch: _pack_ctx(bases, pads, self.epochs.data_ptr(), self.error.data_ptr(), 0, world, ch, 1,
                                   0, int(timeout_s * 1e9))
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

    def allreduce_push(self, offset: int, nbytes: int, dtype: torch.dtype, stage_off: int,
                       scale: Optional[float] = None, channel: int = 0) -> None:
        native._call("b200_allreduce_push", self._ctx[channel], ctypes.c_size_t(offset), ctypes.c_size_t(nbytes),
                     ctypes.c_int(_DTYPE_CODE[dtype]), ctypes.c_float(1.0 / self.world if scale is None else scale),
                     ctypes.c_size_t(stage_off), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def allreduce_wire16(self, offset: int, nbytes: int, wire: torch.dtype, stage_off: int,
                         scale: Optional[float] = None, channel: int = 0) -> None:
        native._call("b200_allreduce_wire16", self._ctx[channel], ctypes.c_size_t(offset), ctypes.c_size_t(nbytes),
                     ctypes.c_int(1 if wire == torch.bfloat16 else 2),
                     ctypes.c_float(1.0 / self.world if scale is None else scale), ctypes.c_size_t(stage_off),
                     ctypes.c_int(self.max_blocks), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def seed_epochs(self, value: int) -> None:
        v = value & 0xFFFFFFFF
        v = v - (1 << 32) if v >= (1 << 31) else v
        self.epochs.fill_(v)
        for p in self.pads:
            p.fill_(v)

    def broadcast(self, offset: int, nbytes: int, root: int, channel: int = 1) -> None:
        native._call("b200_broadcast", self._ctx[channel], ctypes.c_size_t(offset), ctypes.c_size_t(nbytes),
                     ctypes.c_int(root), ctypes.c_int(self.max_blocks),
                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def barrier(self, channel: int = 1) -> None:
        native._call("b200_barrier", self._ctx[channel], ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def error_word(self) -> int:
        return int(self.error[0].item()) & 0xFFFFFFFF


class FusedGradReduce:
    """Weight-gradient GEMM whose epilogue is the all-reduce.  Two forms, both opt-in:

    ``wgrad_nvls``  persistent 2-CTA tcgen05 GEMM; partial tiles go to this rank's copy of ``out``, the kernel's tail
                    all-reduces them in the NVLink switch (multimem), tile by tile, replicas bit-identical.
    ``wgrad``       (round 1) partial tiles are pushed over NVLink to their owner rank, the owner sums them in rank order
                    and pushes the finished tile into every rank's copy of ``out``.

    ``out`` must live in the symmetric arena (e.g. a gradient-bucket slot).  Measured on B200 both are slower than
    GEMM + stand-alone all-reduce, so no model uses them by default.
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

    # -- NVLS form: persistent 2-CTA GEMM whose tail reduces the output through the switch ------------------------
    def _nvls_reserve(self, n_out: int, k_in: int):
        lib = _lib()
        arena = self.comm.arena
        words = int(lib.b200_gemm_tc_nvls_flag_words(ctypes.c_int(n_out), ctypes.c_int(k_in)))
        st = getattr(self, "_nvls_state", None)
        if st is None or st["words"] < words:
            off = arena.alloc(words * 4)
            arena.memory[off:off + words * 4].zero_()
            ep = st["epoch"] if st is not None else None
            if ep is None:
                ep = torch.zeros(4, dtype=torch.int32, device=self.comm.device)
            st = self._nvls_state = {"off": off, "words": words, "epoch": ep}
            torch.cuda.synchronize(self.comm.device)
            dist.barrier(group=self.comm.group)          # every rank's flag object is zeroed before anyone signals
        return st

    def wgrad_nvls(self, dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor) -> bool:
        """``out = mean_r(dy_r^T @ x_r)`` by ONE kernel: the persistent tcgen05 GEMM writes each rank's partial tiles
        into its copy of ``out`` and its tail phase all-reduces them in the NVLink switch (multimem.ld_reduce / st),
        tile by tile as they complete.  Returns False when the shape does not run on that kernel (the caller then runs
        GEMM + all-reduce).  The first call for a larger flag object is a collective (it barriers once)."""
        arena = self.comm.arena
        off = arena.offset_of(out) if arena is not None else None
        if off is None or not arena.multicast or not out.is_contiguous():
            return False
        m_tok, n_out = dy.shape
        k_in = x.shape[1]
        lib = _lib()
        st = self._nvls_reserve(n_out, k_in)
        final_off = int(lib.b200_gemm_tc_nvls_final_off(ctypes.c_int(n_out), ctypes.c_int(k_in)))
        world, rank = self.comm.world_size, self.comm.rank
        pad = [0] * (8 - world)
        raw = struct.pack("<Q8QQQiifi", arena.multicast + off, *[b + st["off"] for b in arena.peer_base], *pad,
                          st["epoch"].data_ptr(), arena.timeout_ns, rank, world, 1.0 / self.comm.divisor, final_off)
        assert len(raw) == lib.b200_nvls_reduce_bytes(), (len(raw), lib.b200_nvls_reduce_bytes())
        nv = ctypes.create_string_buffer(raw, len(raw))
        flags = int(dy.dtype == torch.bfloat16) | (int(out.dtype == torch.bfloat16) << 1)
        desc = (ctypes.c_longlong * 3)(n_out, k_in, k_in)
        err = native.scratch(dy.device).error
        rc = lib.b200_gemm_tc_nvls(ctypes.c_void_p(dy.data_ptr()), ctypes.c_void_p(x.data_ptr()),
                                   ctypes.c_void_p(out.data_ptr()), ctypes.c_int(n_out), ctypes.c_int(k_in),
                                   ctypes.c_int(m_tok), ctypes.c_int(2), ctypes.c_int(flags), desc, nv,
                                   ctypes.c_void_p(err.data_ptr()),
                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc == 801:
            return False
        if rc != 0:
            raise RuntimeError(f"b200_gemm_tc_nvls failed: CUDA error {rc}")
        native._launches += 1
        return True

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
