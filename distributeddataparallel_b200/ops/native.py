"""Bindings + autograd Functions for the sm_100a kernel library (``libb200ddp_kernels.so``).

The library exports a plain C ABI (``include/b200ddp.h``); it is loaded with ctypes and
driven with raw device pointers on the current torch CUDA stream, so every launch is
capturable in a CUDA graph.  If the library is missing on a GPU machine it is built in-tree
with nvcc (sm_100a); failure to load is an error, never a silent fallback.

Gradient placement: a parameter wrapped by our ``DistributedDataParallel`` carries
``param._grad_slot`` -- its strided view inside the flat gradient bucket.  Weight-gradient
kernels write **directly into that slot** and return it as the gradient, so autograd's
``AccumulateGrad`` adopts the slot as ``.grad`` without a copy and the reducer finds the
gradient already in place (SURVEY.md K18/K19 eliminated).
"""
from __future__ import annotations

import ctypes
import os
import struct
from typing import Dict, List, Optional, Sequence

import torch

_lib = None
_launches = 0
_launches_per_capture = 0


# ---------------------------------------------------------------------------------------
# library loading
# ---------------------------------------------------------------------------------------
def load_library():
    global _lib
    if _lib is not None:
        return _lib
    from .. import _build

    path = _build.build_kernels()
    lib = ctypes.CDLL(str(path))
    for name in ("b200_optim_chunk_elems", "b200_optim_rec_bytes"):
        getattr(lib, name).restype = ctypes.c_int
    if "B200DDP_2CTA" in os.environ:       # 0 off, 1 one tile per cluster, 2 persistent (library default)
        lib.b200_gemm_tc_use_2cta(ctypes.c_int(int(os.environ["B200DDP_2CTA"])))
    if "B200DDP_S2_DGRAD" in os.environ:
        lib.b200_gemm_tc_use_s2_dgrad(ctypes.c_int(int(os.environ["B200DDP_S2_DGRAD"])))
    if "B200DDP_FILTER3D" in os.environ:
        lib.b200_gemm_tc_use_filter3d(ctypes.c_int(int(os.environ["B200DDP_FILTER3D"])))
    if "B200DDP_MN32_TMA" in os.environ:
        lib.b200_gemm_tc_use_mn32_tma(ctypes.c_int(int(os.environ["B200DDP_MN32_TMA"])))
    if "B200DDP_IM2COL_TMA" in os.environ:
        lib.b200_gemm_tc_use_im2col_tma(ctypes.c_int(int(os.environ["B200DDP_IM2COL_TMA"])))
    if "B200DDP_TMA" in os.environ:
        lib.b200_gemm_tc_use_tma(ctypes.c_int(int(os.environ["B200DDP_TMA"])))
    _lib = lib
    return lib


def launch_count() -> int:
    return _launches


def launches_per_capture() -> int:
    return _launches_per_capture


def set_launches_per_capture(n: int) -> None:
    global _launches_per_capture
    _launches_per_capture = int(n)


def _call(name: str, *args, kernels: int = 1) -> None:
    global _launches
    lib = load_library()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed: CUDA error {rc} ({torch.cuda.get_device_name()})")
    _launches += kernels


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


_i = ctypes.c_int
_f = ctypes.c_float
_sz = ctypes.c_size_t


# ---------------------------------------------------------------------------------------
# per-device scratch (split-K workspace + tile tickets)
# ---------------------------------------------------------------------------------------
class _Scratch:
    WORKSPACE_BYTES = 32 << 20
    MAX_COUNTERS = 8192

    def __init__(self, device: torch.device) -> None:
        self.workspace = torch.empty(self.WORKSPACE_BYTES // 4, dtype=torch.float32, device=device)
        self.counters = torch.zeros(self.MAX_COUNTERS, dtype=torch.int32, device=device)
        self.row_loss = torch.empty(1 << 16, dtype=torch.float32, device=device)
        self.error = torch.zeros(16, dtype=torch.int32, device=device)     # tcgen05 wait-timeout word
        # zero-on-entry / zero-on-exit accumulators of the one-launch reductions (LayerNorm backward, bias gradients)
        self.red_acc = torch.zeros(32768, dtype=torch.float32, device=device)
        self.red_ticket = torch.zeros(4, dtype=torch.int32, device=device)


_scratch: Dict[tuple, _Scratch] = {}
_lane = 0          # 0: the caller's stream; 1..: a weight-gradient side stream (each lane has its own split-K scratch)


def scratch(device: torch.device) -> _Scratch:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    s = _scratch.get((idx, _lane))
    if s is None:
        s = _scratch[(idx, _lane)] = _Scratch(torch.device("cuda", idx))
    return s


def _ws_args(dev: torch.device):
    s = scratch(dev)
    return (_ptr(s.workspace), _sz(s.WORKSPACE_BYTES), _ptr(s.counters), _i(s.MAX_COUNTERS))


# ---------------------------------------------------------------------------------------
# weight-gradient side streams
# ---------------------------------------------------------------------------------------
# Nothing downstream in the backward pass consumes a weight gradient: only the bucket reduction and the
# optimizer do.  The deep layers of the dpp.py model are latency-bound kernels that leave most SMs idle, so
# wgrad kernels are issued on side streams and run UNDER the dgrad / BatchNorm-backward chain of the following
# layers (the fork/join is captured in the step graph like any other cross-stream edge).  Ordering:
#   * fork: the side stream waits for an event recorded on the caller's stream (dY and x are ready);
#   * every consumer joins: the reducer's bucket launch (parallel/p2p.py, comm.py), the end-of-backward
#     callback queued on the autograd engine, and the fused optimizers.
# Tensors read on a side stream are ``record_stream``-ed so the caching allocator does not recycle them early.
_SIDE_STREAMS = int(os.environ.get("B200DDP_WGRAD_STREAMS", "1"))
_side: Dict[int, list] = {}          # device index -> [streams]
_side_next: Dict[int, int] = {}
_side_dirty: Dict[int, set] = {}     # device index -> lanes with work the caller's stream has not joined yet
_join_queued = False
_join_task = None          # autograd graph-task id the end-of-backward join was queued for


def set_wgrad_streams(n: int) -> None:
    """0 disables the side streams (weight gradients run in line on the caller's stream)."""
    global _SIDE_STREAMS
    _SIDE_STREAMS = max(0, int(n))


def _side_stream(dev: torch.device):
    """(lane, stream) of the next weight-gradient side stream, or (0, None) when disabled."""
    if _SIDE_STREAMS <= 0:
        return 0, None
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    pool = _side.get(idx)
    if pool is None or len(pool) != _SIDE_STREAMS:
        pool = _side[idx] = [torch.cuda.Stream(device=idx) for _ in range(_SIDE_STREAMS)]
    k = _side_next.get(idx, 0) % len(pool)
    _side_next[idx] = k + 1
    return k + 1, pool[k]


def side_stream_events(device=None):
    """One fresh event per side stream that has un-joined work: whoever consumes weight gradients on another
    stream (the bucket all-reduce on the communication stream) waits on these."""
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    out = []
    for lane in sorted(_side_dirty.get(idx, ())):
        ev = torch.cuda.Event()
        ev.record(_side[idx][lane - 1])
        out.append(ev)
    return out


def join_side_streams(device=None) -> None:
    """Order the current stream after all outstanding side-stream work."""
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    cur = torch.cuda.current_stream(idx)
    for ev in side_stream_events(idx):
        cur.wait_event(ev)
    _side_dirty.pop(idx, None)


def _graph_task_id():
    fn = getattr(torch._C, "_current_graph_task_id", None)
    return fn() if fn is not None else None


def _end_of_backward_join() -> None:
    global _join_queued
    _join_queued = False
    for idx in list(_side_dirty):
        join_side_streams(idx)


def _run_on_side(dev: torch.device, tensors, fn) -> None:
    """Run ``fn()`` (kernel launches) on a side stream forked from the current one."""
    global _lane, _join_queued, _join_task
    lane, stream = _side_stream(dev)
    if stream is None:
        fn()
        return
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(idx))
    stream.wait_event(ev)
    prev = _lane
    _lane = lane
    try:
        with torch.cuda.stream(stream):
            fn()
    finally:
        _lane = prev
    for t in tensors:
        if t is not None:
            t.record_stream(stream)
    _side_dirty.setdefault(idx, set()).add(lane)
    task = _graph_task_id()
    if not _join_queued or task != _join_task:   # (a backward pass that raised never ran its callbacks: re-queue)
        try:
            torch.autograd.Variable._execution_engine.queue_callback(_end_of_backward_join)
            _join_queued, _join_task = True, task
        except RuntimeError:
            join_side_streams(idx)       # not inside a backward pass: join right away


# ---------------------------------------------------------------------------------------
# tensor-core (tcgen05) path switch
# ---------------------------------------------------------------------------------------
import os as _os

_tc_mode = _os.environ.get("B200DDP_TC", "auto")       # auto | 1 (always tcgen05) | 0 (always FFMA)
_NOT_SUPPORTED = 801          # cudaErrorNotSupported: shape needs the FFMA fallback
tc_launches = 0
ffma_launches = 0
_choice: Dict[tuple, int] = {}        # (entry point, geometry) -> 0: FFMA kernel, else split-K cap of the tcgen05 kernel


def set_tensor_cores(enabled) -> None:
    """True: always the tcgen05 kernels; False: always the FFMA kernels; "auto": time both the
    first time a shape is seen (outside graph capture) and keep the faster one."""
    global _tc_mode
    _tc_mode = "auto" if enabled == "auto" else ("1" if enabled else "0")


def tensor_cores_enabled() -> bool:
    return _tc_mode != "0"


def conv_choices() -> Dict[tuple, int]:
    return dict(_choice)


# ---- tuned choices: persisted per GPU model, identical on every rank ------------------------------------------------
# The table maps (entry point, geometry) -> kernel variant.  It is loaded from ``tuning/<gpu name>.json`` next to the
# package (shipped with the repository for B200, extended on the fly for new shapes), so tests, ``smoke()`` and
# training runs do not re-time known shapes, and every rank of a job runs the same variants -- per-rank timing noise
# can no longer make ranks round differently.  ``B200DDP_TUNE=retune`` ignores the file; ``=off`` never writes it.
_TUNE_MODE = os.environ.get("B200DDP_TUNE", "auto")
_tune_loaded = False
_tune_dirty = False


def _tune_path():
    from pathlib import Path

    name = "".join(ch if ch.isalnum() else "_" for ch in torch.cuda.get_device_name())
    return Path(__file__).resolve().parent.parent / "tuning" / f"{name}.json"


def _key_str(key: tuple) -> str:
    return key[0] + "|" + ",".join(str(v) for v in key[1:])


def _key_from(s: str) -> tuple:
    name, rest = s.split("|", 1)
    return (name,) + tuple(int(v) for v in rest.split(",") if v != "")


def _load_choices() -> None:
    global _tune_loaded
    if _tune_loaded:
        return
    _tune_loaded = True
    if _TUNE_MODE == "retune" or os.environ.get("B200DDP_TC_VARIANTS"):
        return
    import json

    try:
        table = json.loads(_tune_path().read_text())
    except (OSError, ValueError):
        return
    for k, v in table.get("choices", {}).items():
        _choice.setdefault(_key_from(k), int(v))


def save_choices(force: bool = False) -> None:
    """Write the table (atomic rename); only local rank 0 writes, and only when something new was tuned."""
    global _tune_dirty
    if not (_tune_dirty or force) or _TUNE_MODE == "off" or os.environ.get("B200DDP_TC_VARIANTS"):
        return
    if int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")) or 0) != 0:
        return
    import json

    path = _tune_path()
    try:
        path.parent.mkdir(parents=True, exist_ok=True)
        body = {"gpu": torch.cuda.get_device_name(), "format": 1,
                "choices": {_key_str(k): v for k, v in sorted(_choice.items(), key=lambda kv: _key_str(kv[0]))}}
        tmp = path.with_suffix(f".tmp{os.getpid()}")
        tmp.write_text(json.dumps(body, indent=0, sort_keys=True))
        os.replace(tmp, path)
        _tune_dirty = False
    except OSError:
        pass


def sync_choices(group=None) -> None:
    """Collective: all ranks end up with the same table -- the union of what any rank tuned, the lowest rank winning
    where two ranks timed the same shape differently."""
    import torch.distributed as dist

    global _tune_dirty
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    tables = [None] * dist.get_world_size(group)
    dist.all_gather_object(tables, {_key_str(k): v for k, v in _choice.items()}, group=group)
    merged = {}
    for t in reversed(tables):          # lowest rank wins on conflicts
        merged.update(t)
    before = dict(_choice)
    _choice.clear()
    _choice.update({_key_from(k): int(v) for k, v in merged.items()})
    if _choice != before or dist.get_rank(group) == 0:
        _tune_dirty = _tune_dirty or _choice != before


import atexit as _atexit

_atexit.register(save_choices)


# split-K variants the tuner tries for a tcgen05 convolution: (upper bound on the split factor, reduce the splits
# through a thread-block cluster's distributed shared memory instead of an L2 workspace); a choice is stored as
# cap + 100 * cluster
_TC_VARIANTS = ((32, 0), (1, 0), (4, 0), (12, 0), (16, 1), (8, 1), (4, 1), (2, 1))
if os.environ.get("B200DDP_TC_VARIANTS"):          # e.g. "32,1,4,108": restrict what the tuner may pick (A/B runs)
    _TC_VARIANTS = tuple((int(v) % 100, int(v) // 100) for v in os.environ["B200DDP_TC_VARIANTS"].split(","))
_FORCE_CLUSTER = os.environ.get("B200DDP_CLUSTER_REDUCE", "") == "1"     # testing: cluster reduction wherever K is split


def _launch_tc(name, head, geom, dev, variant: int = 32) -> int:
    s = scratch(dev)
    lib = load_library()
    cap, cluster = variant % 100, variant // 100
    if _FORCE_CLUSTER and variant == 32:
        cap, cluster = 8, 1
    if cap == 32 and not cluster:
        return getattr(lib, name + "_tc")(*head, *geom, *_ws_args(dev), _ptr(s.error), _stream())
    lib.b200_gemm_tc_split_cap(_i(cap))            # library-wide knobs: restore them for every other GEMM caller
    lib.b200_gemm_tc_cluster_reduce(_i(cluster))
    try:
        return getattr(lib, name + "_tc")(*head, *geom, *_ws_args(dev), _ptr(s.error), _stream())
    finally:
        lib.b200_gemm_tc_split_cap(_i(32))
        lib.b200_gemm_tc_cluster_reduce(_i(0))


def _launch_ffma(name, head, geom, dev) -> int:
    lib = load_library()
    if name == "b200_conv2d_fprop":
        x, w, y, stats = head
        return lib.b200_conv2d_fprop(x, w, _ptr(None), y, stats, *geom, *_ws_args(dev), _stream())
    return getattr(lib, name)(*head, *geom, *_ws_args(dev), _stream())


_flush_buf: Dict[int, torch.Tensor] = {}


def _time_variant(fn, acts=(), iters: int = 6) -> float:
    """Device time of one call under the cache state it meets inside a training step: before every launch the L2 is
    flushed (a 160 MB memset, larger than the 126 MB L2) and the activation operands ``acts`` -- which in a real
    step were just written by the previous kernel -- are read back in, so filters come from HBM and activations
    from L2.  The launches are captured in a CUDA graph and replayed (no host launch cost in the measurement); the
    flush and re-warm cost is the same for every variant, so totals are compared directly."""
    fn()
    torch.cuda.synchronize()
    dev = torch.cuda.current_device()
    buf = _flush_buf.get(dev)
    if buf is None:
        buf = _flush_buf[dev] = torch.empty(160 << 20, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
        for _ in range(iters):
            buf.zero_()
            for t in acts:
                t.sum()
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    g.replay()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / (2 * iters)


def _conv_call(name: str, head: tuple, geom: tuple, dev: torch.device, acts=()) -> None:
    """Dispatch one convolution pass to the tcgen05 kernel (with a tuned split-K bound) or to the FFMA kernel."""
    global _launches, tc_launches, ffma_launches, _tune_dirty
    choice = 32 if _tc_mode == "1" else 0          # 0 = FFMA, otherwise the split-K cap of the tcgen05 call
    if _tc_mode == "auto":
        _load_choices()
        key = (name,) + tuple(g.value for g in geom)
        choice = _choice.get(key)
        if choice is None:
            if torch.cuda.is_current_stream_capturing():
                choice = 0                           # cannot time inside a capture: stay on the safe path
            else:
                # time the variants on the real operands; statistics / accumulation are switched off while
                # timing so the repeated runs have no side effects beyond rewriting the output
                quiet = head[:3] + (_ptr(None),) if name == "b200_conv2d_fprop" else head
                if name == "b200_conv2d_wgrad":
                    quiet = head[:3] + (_i(0),)
                choice = 0
                if _launch_tc(name, quiet, geom, dev) == 0:
                    best = _time_variant(lambda: _launch_ffma(name, quiet, geom, dev), acts)
                    for cap, cluster in _TC_VARIANTS:
                        v = cap + 100 * cluster
                        t = _time_variant(lambda: _launch_tc(name, quiet, geom, dev, v), acts)
                        if t < best:
                            best, choice = t, v
                _choice[key] = choice
                _tune_dirty = True
    if choice:
        rc = _launch_tc(name, head, geom, dev, choice)
        if rc == 0:
            _launches += 1
            tc_launches += 1
            return
        if rc != _NOT_SUPPORTED:
            raise RuntimeError(f"{name}_tc failed: CUDA error {rc}")
    rc = _launch_ffma(name, head, geom, dev)
    if rc != 0:
        raise RuntimeError(f"{name} failed: CUDA error {rc}")
    _launches += 1
    ffma_launches += 1


def stamp(ptr: int) -> None:
    """Write %globaltimer (ns) to the 64-bit word at device address ``ptr`` on the current stream."""
    _call("b200_stamp", ctypes.c_void_p(ptr), _stream())


def tc_error_word(device=None) -> int:
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    word = 0
    for (d, _lane_id), s in list(_scratch.items()):          # every lane (caller's stream + side streams)
        if d == idx:
            word |= int(s.error[0].item()) & 0xFFFFFFFF
    return word


def gemm_tc(a: torch.Tensor, b: torch.Tensor, mode: int, bias: Optional[torch.Tensor] = None,
            out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """Dense tcgen05 GEMM, fp32 output.  mode 0: a[M,K] b[N,K]^T; 1: a[M,K] b[K,N]; 2: a[K,M]^T b[K,N]."""
    assert a.dtype == b.dtype and a.dtype in (torch.float32, torch.bfloat16) and a.is_contiguous() and b.is_contiguous()
    if mode == 0:
        (m, k), n = a.shape, b.shape[0]
    elif mode == 1:
        (m, k), n = a.shape, b.shape[1]
    else:
        (k, m), n = a.shape, b.shape[1]
    if out is None:
        out = torch.empty((m, n), dtype=torch.float32, device=a.device)
    s = scratch(a.device)
    _call("b200_gemm_tc", _ptr(a), _ptr(b), _ptr(out), _ptr(bias), _i(m), _i(n), _i(k), _i(mode),
          _i(int(a.dtype == torch.bfloat16)), _i(int(accumulate)), *_ws_args(a.device), _ptr(s.error), _stream())
    return out


# ---------------------------------------------------------------------------------------
# layout helpers
# ---------------------------------------------------------------------------------------
def _nhwc(x: torch.Tensor) -> torch.Tensor:
    """4-D tensor, logical NCHW, physically dense NHWC, fp32."""
    if x.dtype != torch.float32:
        x = x.float()
    return x.contiguous(memory_format=torch.channels_last)


def _empty_nhwc(n, c, h, w, device) -> torch.Tensor:
    return torch.empty((n, c, h, w), dtype=torch.float32, device=device, memory_format=torch.channels_last)


def _grad_target(param: torch.Tensor, slot: Optional[torch.Tensor], like: torch.Tensor):
    """(tensor to write the gradient into, accumulate?)  Prefer the DDP bucket slot."""
    # A parameter used by two ops in one forward (a shared layer) reaches here twice with ``grad is None`` --
    # AccumulateGrad only runs after both contributions exist -- so the slot is handed out once per backward
    # pass (the DDP wrapper clears the claim in forward); later uses get private memory and autograd adds them up.
    if slot is not None and param.grad is None and slot.shape == like.shape and not getattr(param, "_slot_claimed", False):
        if hasattr(param, "_ddp_index"):
            param._slot_claimed = True
        return slot.detach(), False          # fresh alias of the slot: AccumulateGrad adopts it
    return torch.empty_like(like, memory_format=torch.preserve_format), False


# ---------------------------------------------------------------------------------------
# input pipeline
# ---------------------------------------------------------------------------------------
def preprocess_u8(x_u8: torch.Tensor, mean: float = 0.5, std: float = 0.5) -> torch.Tensor:
    assert x_u8.dtype == torch.uint8 and x_u8.dim() == 4, "expected uint8 [N,H,W,C]"
    x_u8 = x_u8.contiguous()
    n, h, w, c = x_u8.shape
    out = torch.empty((n, h, w, c), dtype=torch.float32, device=x_u8.device)
    _call("b200_preprocess_u8", _ptr(x_u8), _ptr(out), _sz(x_u8.numel()), _f(mean), _f(std), _stream())
    return out.permute(0, 3, 1, 2)       # logical NCHW view of NHWC storage (== channels_last)


# ---------------------------------------------------------------------------------------
# convolution
# ---------------------------------------------------------------------------------------
def _conv_geom(x: torch.Tensor, w: torch.Tensor, stride: int, pad: int):
    n, c, h, wd = x.shape
    k, c2, r, s = w.shape
    assert c == c2, f"channel mismatch {c} vs {c2}"
    p = (h + 2 * pad - r) // stride + 1
    q = (wd + 2 * pad - s) // stride + 1
    return n, h, wd, c, k, r, s, p, q


def _pad_channels(t: torch.Tensor, c_to: int) -> torch.Tensor:
    """[N, C, H, W] channels-last -> [N, c_to, H, W] channels-last with zero-filled extra channels."""
    n, c, h, w = t.shape
    out = t.new_zeros((n, h, w, c_to))
    out[..., :c] = t.permute(0, 2, 3, 1)
    return out.permute(0, 3, 1, 2)


class _Conv2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, pad, stats):
        xc, wc = _nhwc(x), _nhwc(w)
        c_real = xc.shape[1]
        # An RGB stem (C = 3) has no 16-byte channel vectors, which both the tcgen05 gathers and the vectorised
        # FFMA loads need: run it as a 4-channel convolution with a zero fourth channel (its filter gradient
        # is exactly zero and is dropped again in backward).
        padded = c_real % 4 != 0 and _tc_mode != "0" and not ctx.needs_input_grad[0]
        if padded:
            c4 = (c_real + 3) // 4 * 4
            xc, wc = _pad_channels(xc, c4), _pad_channels(wc, c4)
        n, h, wd, c, k, r, s, p, q = _conv_geom(xc, wc, stride, pad)
        y = _empty_nhwc(n, k, p, q, x.device)
        _conv_call("b200_conv2d_fprop", (_ptr(xc), _ptr(wc), _ptr(y), _ptr(stats)),
                   (_i(n), _i(h), _i(wd), _i(c), _i(k), _i(r), _i(s), _i(stride), _i(pad)), x.device, acts=(xc,))
        ctx.save_for_backward(xc, wc)
        ctx.geom = (n, h, wd, c, k, r, s, stride, pad)
        ctx.param = w
        ctx.slot = getattr(w, "_grad_slot", None)
        ctx.c_real = c_real if padded else None
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, wc = ctx.saved_tensors
        n, h, wd, c, k, r, s, stride, pad = ctx.geom
        dy = _nhwc(dy)
        geom = (_i(n), _i(h), _i(wd), _i(c), _i(k), _i(r), _i(s), _i(stride), _i(pad))
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = _empty_nhwc(n, c, h, wd, dy.device)
            _conv_call("b200_conv2d_dgrad", (_ptr(dy), _ptr(wc), _ptr(dx)), geom, dy.device, acts=(dy,))
        if ctx.needs_input_grad[1]:
            if ctx.c_real is None:
                dw, acc = _grad_target(ctx.param, ctx.slot, wc)
                key = ("b200_conv2d_wgrad",) + tuple(g.value for g in geom)
                _load_choices()
                if _tc_mode == "auto" and key not in _choice and not torch.cuda.is_current_stream_capturing():
                    # first sight of the shape: tune in line (the tuner synchronises), side streams from the next call
                    _conv_call("b200_conv2d_wgrad", (_ptr(dy), _ptr(xc), _ptr(dw), _i(int(acc))), geom, dy.device, acts=(dy, xc))
                else:
                    _run_on_side(dy.device, (dy, xc, dw), lambda: _conv_call(
                        "b200_conv2d_wgrad", (_ptr(dy), _ptr(xc), _ptr(dw), _i(int(acc))), geom, dy.device, acts=(dy, xc)))
            else:
                dw_pad = torch.empty_like(wc)
                _conv_call("b200_conv2d_wgrad", (_ptr(dy), _ptr(xc), _ptr(dw_pad), _i(0)), geom, dy.device, acts=(dy, xc))
                dw, acc = _grad_target(ctx.param, ctx.slot, _nhwc(ctx.param))
                if acc:
                    dw.add_(dw_pad[:, :ctx.c_real])
                else:
                    dw.copy_(dw_pad[:, :ctx.c_real])
        return dx, dw, None, None, None


def conv2d(x, w, stride: int, padding: int, stats: Optional[torch.Tensor] = None):
    return _Conv2d.apply(x, w, int(stride), int(padding), stats)


# ---------------------------------------------------------------------------------------
# batch norm (+ residual) (+ relu)
# ---------------------------------------------------------------------------------------
_FUSED_BN_BWD = _os.environ.get("B200DDP_FUSED_BN_BWD", "0") != "0"   # measured: no gain (step is not launch bound)
# Cluster BatchNorm kernels (one launch per direction, deterministic) for tensors of <= 131 072 elements:
#   backward: on by default (one kernel instead of reduce + apply on the backward critical path);
#   forward : B200DDP_BN_STATS=exact computes exact two-pass statistics in the BatchNorm kernel itself instead of taking
#             the float-atomic sums of the convolution epilogue (bit-reproducible, robust for |mean| >> std).
_DETERMINISTIC = _os.environ.get("B200DDP_DETERMINISTIC", "0") != "0"
_BN_CLUSTER_BWD = _os.environ.get("B200DDP_BN_CLUSTER", "1" if _DETERMINISTIC else "0") != "0"
_BN_STATS_MODE = _os.environ.get("B200DDP_BN_STATS", "exact" if _DETERMINISTIC else "fused")          # fused | exact


def set_bn_stats_mode(mode: str) -> None:
    global _BN_STATS_MODE
    assert mode in ("fused", "exact")
    _BN_STATS_MODE = mode


def bn_stats_mode() -> str:
    return _BN_STATS_MODE


def _cluster_bn_ok(m: int, c: int) -> bool:
    return c % 4 == 0 and m * c <= 131072 and (c // 4) <= 512 and 512 % (c // 4) == 0


class BNWorkspace:
    """Per-layer statistic accumulators.  ``fwd`` is filled by the producing conv's epilogue
    (or by bn_stats), ``bwd`` by bn_bwd_reduce; each apply kernel zeroes the *other* buffer so
    steady-state training needs no memset launches."""

    def __init__(self, channels: int, device) -> None:
        self.fwd = torch.zeros(2 * channels, dtype=torch.float32, device=device)
        self.bwd = torch.zeros(2 * channels, dtype=torch.float32, device=device)
        self.tickets = torch.zeros(2, dtype=torch.int32, device=device)      # grid barrier of the fused backward
        self.fwd_dirty = False
        self.bwd_dirty = False

    def begin_forward(self) -> None:
        if self.fwd_dirty:           # previous forward was never followed by a backward
            self.fwd.zero_()
        self.fwd_dirty = True


def bn_workspace(module, channels: int, device) -> BNWorkspace:
    ws = module.__dict__.get("_b200_ws")
    if ws is None or ws.fwd.device != device or ws.fwd.numel() != 2 * channels:
        ws = BNWorkspace(channels, device)
        module.__dict__["_b200_ws"] = ws
    return ws


class _BatchNormAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, rmean, rvar, nbt, residual, relu, training, momentum, eps, ws, stats_ready):
        xc = _nhwc(x)
        n, c, h, w = xc.shape
        m = n * h * w
        res = _nhwc(residual) if residual is not None else None
        out = _empty_nhwc(n, c, h, w, x.device)
        save_mean = torch.empty(c, dtype=torch.float32, device=x.device)
        save_invstd = torch.empty(c, dtype=torch.float32, device=x.device)
        if training and not stats_ready and _BN_STATS_MODE == "exact" and _cluster_bn_ok(m, c):
            _call("b200_bn_fwd_cluster", _ptr(xc), _ptr(gamma), _ptr(beta), _ptr(res), _ptr(out), _ptr(rmean), _ptr(rvar),
                  _ptr(nbt), _ptr(save_mean), _ptr(save_invstd), _ptr(ws.bwd), _i(ws.bwd.numel()), _i(m), _i(c),
                  _f(momentum), _f(eps), _i(int(relu)), _stream())
        else:
            if training and not stats_ready:
                _call("b200_bn_stats", _ptr(xc), _ptr(ws.fwd), _i(m), _i(c), _stream())
            _call("b200_bn_fwd_apply", _ptr(xc), _ptr(ws.fwd), _ptr(gamma), _ptr(beta), _ptr(res), _ptr(out),
                  _ptr(rmean), _ptr(rvar), _ptr(nbt), _ptr(save_mean), _ptr(save_invstd),
                  _ptr(ws.bwd), _i(ws.bwd.numel()), _i(m), _i(c), _f(momentum), _f(eps), _i(int(relu)),
                  _i(int(training)), _stream())
        ws.bwd_dirty = False
        ctx.save_for_backward(xc, out if relu else None, gamma, save_mean, save_invstd)
        ctx.cfg = (m, c, bool(relu), residual is not None, training)
        ctx.ws = ws
        ctx.params = (gamma, beta)
        ctx.slots = (getattr(gamma, "_grad_slot", None), getattr(beta, "_grad_slot", None))
        if not training:
            # eval: normalisation uses running stats; backward treats them as constants
            pass
        return out

    @staticmethod
    def backward(ctx, dout):
        xc, out, gamma, save_mean, save_invstd = ctx.saved_tensors
        m, c, relu, has_res, training = ctx.cfg
        ws = ctx.ws
        dout = _nhwc(dout)
        if not training:
            raise RuntimeError("backward through eval-mode BatchNorm is not supported by the native kernels")
        if ws.bwd_dirty:
            ws.bwd.zero_()
        ws.bwd_dirty = True
        dx = torch.empty_like(xc, memory_format=torch.preserve_format)
        dres = torch.empty_like(xc, memory_format=torch.preserve_format) if has_res else None
        dgamma, _ = _grad_target(ctx.params[0], ctx.slots[0], gamma)
        dbeta, _ = _grad_target(ctx.params[1], ctx.slots[1], gamma)
        if _BN_CLUSTER_BWD and _cluster_bn_ok(m, c):
            _call("b200_bn_bwd_cluster", _ptr(dout), _ptr(out), _ptr(xc), _ptr(gamma), _ptr(save_mean), _ptr(save_invstd),
                  _ptr(dx), _ptr(dres), _ptr(dgamma), _ptr(dbeta), _i(0), _ptr(ws.fwd), _i(ws.fwd.numel()), _i(m), _i(c),
                  _i(int(relu)), _stream())
        elif _FUSED_BN_BWD:
            _call("b200_bn_bwd_fused", _ptr(dout), _ptr(out), _ptr(xc), _ptr(gamma), _ptr(save_mean),
                  _ptr(save_invstd), _ptr(ws.bwd), _ptr(dx), _ptr(dres), _ptr(dgamma), _ptr(dbeta), _i(0),
                  _ptr(ws.fwd), _i(ws.fwd.numel()), _ptr(ws.tickets), _i(m), _i(c), _i(int(relu)), _stream())
        else:
            _call("b200_bn_bwd_reduce", _ptr(dout), _ptr(out), _ptr(xc), _ptr(save_mean), _ptr(save_invstd),
                  _ptr(ws.bwd), _i(m), _i(c), _i(int(relu)), _stream())
            _call("b200_bn_bwd_apply", _ptr(dout), _ptr(out), _ptr(xc), _ptr(gamma), _ptr(save_mean),
                  _ptr(save_invstd), _ptr(ws.bwd), _ptr(dx), _ptr(dres), _ptr(dgamma), _ptr(dbeta), _i(0),
                  _ptr(ws.fwd), _i(ws.fwd.numel()), _i(m), _i(c), _i(int(relu)), _stream())
        ws.fwd_dirty = False
        return dx, dgamma, dbeta, None, None, None, dres, None, None, None, None, None, None


def batch_norm_act(x, weight, bias, running_mean, running_var, num_batches_tracked, residual, relu,
                   training, momentum, eps, ws: Optional[BNWorkspace] = None, stats_ready: bool = False):
    if ws is None:
        ws = BNWorkspace(x.shape[1], x.device)
        stats_ready = False
    elif not stats_ready and training:
        ws.begin_forward()
    return _BatchNormAct.apply(x, weight, bias, running_mean, running_var, num_batches_tracked, residual,
                               bool(relu), bool(training), float(momentum), float(eps), ws, bool(stats_ready))


def conv_bn_act(x, conv_weight, stride, padding, bn, residual=None, relu=False):
    """conv -> BN(+residual)(+ReLU) with the BN statistics accumulated by the conv epilogue."""
    training = bn.training or not bn.track_running_stats
    ws = bn_workspace(bn, conv_weight.shape[0], x.device)
    if training and _BN_STATS_MODE == "exact":
        # statistics are computed by the BatchNorm kernel itself (exact two-pass, deterministic): the convolution
        # epilogue accumulates nothing
        y = conv2d(x, conv_weight, stride, padding, None)
        return batch_norm_act(
            y, bn.weight, bn.bias, bn.running_mean if bn.track_running_stats else None,
            bn.running_var if bn.track_running_stats else None,
            bn.num_batches_tracked if (bn.training and bn.track_running_stats) else None,
            residual, relu, training, 0.1 if bn.momentum is None else bn.momentum, bn.eps, ws, stats_ready=False)
    if training:
        ws.begin_forward()
    y = conv2d(x, conv_weight, stride, padding, ws.fwd if training else None)
    return batch_norm_act(
        y, bn.weight, bn.bias, bn.running_mean if bn.track_running_stats else None,
        bn.running_var if bn.track_running_stats else None,
        bn.num_batches_tracked if (bn.training and bn.track_running_stats) else None,
        residual, relu, training, 0.1 if bn.momentum is None else bn.momentum, bn.eps, ws, stats_ready=training)


# ---------------------------------------------------------------------------------------
# pooling
# ---------------------------------------------------------------------------------------
class _MaxPool3x3s2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xc = _nhwc(x)
        n, c, h, w = xc.shape
        p, q = (h - 1) // 2 + 1, (w - 1) // 2 + 1
        y = _empty_nhwc(n, c, p, q, x.device)
        idx = torch.empty((n, p, q, c), dtype=torch.uint8, device=x.device)
        _call("b200_maxpool3x3s2_fwd", _ptr(xc), _ptr(y), _ptr(idx), _i(n), _i(h), _i(w), _i(c), _stream())
        ctx.save_for_backward(idx)
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.shape
        dy = _nhwc(dy)
        dx = _empty_nhwc(n, c, h, w, dy.device)
        _call("b200_maxpool3x3s2_bwd", _ptr(dy), _ptr(idx), _ptr(dx), _i(n), _i(h), _i(w), _i(c), _stream())
        return dx


def max_pool_3x3s2(x):
    return _MaxPool3x3s2.apply(x)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xc = _nhwc(x)
        n, c, h, w = xc.shape
        y = torch.empty((n, c), dtype=torch.float32, device=x.device)
        _call("b200_avgpool_fwd", _ptr(xc), _ptr(y), _i(n), _i(h * w), _i(c), _stream())
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        n, c, h, w = ctx.shape
        dy = dy.contiguous()
        dx = _empty_nhwc(n, c, h, w, dy.device)
        _call("b200_avgpool_bwd", _ptr(dy), _ptr(dx), _i(n), _i(h * w), _i(c), _stream())
        return dx


def global_avg_pool(x):
    n, c, h, w = x.shape
    if h * w == 1:
        return _nhwc(x).reshape(n, c)      # identity for 1x1 maps (the dpp.py case): a view, no kernel
    return _GlobalAvgPool.apply(x)


# ---------------------------------------------------------------------------------------
# linear = 1x1 convolution over a [M,1,1,K] tensor
# ---------------------------------------------------------------------------------------
class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.reshape(-1, x.shape[-1]).contiguous().float()
        wc = w.contiguous()
        m, k = x2.shape
        n = wc.shape[0]
        y = torch.empty((m, n), dtype=torch.float32, device=x.device)
        _call("b200_conv2d_fprop", _ptr(x2), _ptr(wc), _ptr(b), _ptr(y), _ptr(None),
              _i(m), _i(1), _i(1), _i(k), _i(n), _i(1), _i(1), _i(1), _i(0), *_ws_args(x.device), _stream())
        ctx.save_for_backward(x2, wc)
        ctx.lead = x.shape[:-1]
        ctx.params = (w, b)
        ctx.slots = (getattr(w, "_grad_slot", None), getattr(b, "_grad_slot", None) if b is not None else None)
        return y.reshape(*ctx.lead, n)

    @staticmethod
    def backward(ctx, dy):
        x2, wc = ctx.saved_tensors
        m, k = x2.shape
        n = wc.shape[0]
        dy2 = dy.reshape(m, n).contiguous()
        geom = (_i(m), _i(1), _i(1), _i(k), _i(n), _i(1), _i(1), _i(1), _i(0))
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((m, k), dtype=torch.float32, device=dy.device)
            _call("b200_conv2d_dgrad", _ptr(dy2), _ptr(wc), _ptr(dx), *geom, *_ws_args(dy.device), _stream())
            dx = dx.reshape(*ctx.lead, k)
        if ctx.needs_input_grad[1]:
            dw, _ = _grad_target(ctx.params[0], ctx.slots[0], wc)
            _call("b200_conv2d_wgrad", _ptr(dy2), _ptr(x2), _ptr(dw), _i(0), *geom, *_ws_args(dy.device), _stream())
        if ctx.params[1] is not None and ctx.needs_input_grad[2]:
            db, _ = _grad_target(ctx.params[1], ctx.slots[1], ctx.params[1])
            _call("b200_colsum", _ptr(dy2), _ptr(db), _i(m), _i(n), _i(0), _stream())
        return dx, dw, db


def linear(x, w, b=None):
    return _Linear.apply(x, w, b)


# ---------------------------------------------------------------------------------------
# loss
# ---------------------------------------------------------------------------------------
class _SoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        z = logits.contiguous().float()
        b, v = z.shape
        loss = torch.empty((), dtype=torch.float32, device=z.device)
        dz = torch.empty_like(z)
        sc = scratch(z.device)
        row = sc.row_loss if b <= sc.row_loss.numel() else torch.empty(b, dtype=torch.float32, device=z.device)
        _call("b200_softmax_xent", _ptr(z), _ptr(target.contiguous()), _ptr(loss), _ptr(dz), _ptr(row), _i(b), _i(v),
              _stream(), kernels=1 if b <= 32 else 2)
        ctx.save_for_backward(dz)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (dz,) = ctx.saved_tensors
        out = torch.empty_like(dz)
        _call("b200_scale", _ptr(dz), _ptr(out), _sz(dz.numel()), _ptr(gout.contiguous()), _f(1.0), _stream())
        return out, None


def cross_entropy(logits, target):
    return _SoftmaxXent.apply(logits, target)


# ---------------------------------------------------------------------------------------
# optimizers
# ---------------------------------------------------------------------------------------
def _is_dense(t: torch.Tensor) -> bool:
    """non-overlapping and dense: some permutation of the dims is contiguous."""
    expect = 1
    for size, stride in sorted(((sz, st) for sz, st in zip(t.shape, t.stride()) if sz != 1), key=lambda p: p[1]):
        if stride != expect:
            return False
        expect *= size
    return True


def _dense_same_layout(ts: Sequence[torch.Tensor]) -> bool:
    """p and g share dtype (fp32 or bf16) and layout; optimizer state tensors are fp32, same layout."""
    s0 = ts[0]
    if ts[0].dtype not in (torch.float32, torch.bfloat16) or ts[1].dtype != ts[0].dtype:
        return False
    return all(_is_dense(t) and t.stride() == s0.stride() for t in ts) and all(t.dtype == torch.float32 for t in ts[2:])


class _OptimTable:
    def __init__(self, ps, gs, s0s, s1s, device) -> None:
        lib = load_library()
        chunk = lib.b200_optim_chunk_elems()
        assert lib.b200_optim_rec_bytes() == 48
        recs = bytearray()
        chunks = bytearray()
        for ti, p in enumerate(ps):
            tensors = [p, gs[ti]] + ([s0s[ti]] if s0s else []) + ([s1s[ti]] if s1s else [])
            if not _dense_same_layout(tensors):
                raise RuntimeError("fused optimizer needs dense fp32/bf16 params+grads (fp32 state) with identical strides")
            recs += struct.pack("<QQQQqq", p.data_ptr(), gs[ti].data_ptr(), s0s[ti].data_ptr() if s0s else 0,
                                s1s[ti].data_ptr() if s1s else 0, p.numel(), int(p.dtype == torch.bfloat16))
            for c in range(-(-p.numel() // chunk)):
                chunks += struct.pack("<ii", ti, c)
        self.num_chunks = len(chunks) // 8
        self.num_tensors = len(ps)
        self.recs = torch.frombuffer(recs, dtype=torch.uint8).to(device)
        self.chunks = torch.frombuffer(chunks, dtype=torch.uint8).to(device)
        self.hyper = torch.zeros(16, dtype=torch.float32, device=device)
        self.hyper_host: Optional[tuple] = None
        self.step_ptrs = None

    def set_hyper(self, vals: Sequence[float], keep_tail: int = 0) -> None:
        vals = tuple(float(v) for v in vals)
        if vals != self.hyper_host:
            n = len(vals)
            self.hyper[:n].copy_(torch.tensor(vals, dtype=torch.float32), non_blocking=False)
            self.hyper_host = vals


def _table(tables: dict, gi: int, ps, gs, s0s, s1s) -> _OptimTable:
    key = (gi, tuple(p.data_ptr() for p in ps), tuple(g.data_ptr() for g in gs),
           tuple(s.data_ptr() for s in (s0s or ())), tuple(s.data_ptr() for s in (s1s or ())))
    tab = tables.get(gi)
    if tab is None or tab[0] != key:
        tab = (key, _OptimTable(ps, gs, s0s, s1s, ps[0].device))
        tables[gi] = tab
    return tab[1]


def fused_sgd(tables: dict, gi: int, ps, gs, bufs, *, lr, momentum, dampening, weight_decay, nesterov, maximize,
              first_step: bool, grad_scale: float = 1.0) -> None:
    tab = _table(tables, gi, ps, gs, bufs, None)
    tab.set_hyper((lr, momentum, dampening, weight_decay, float(nesterov), float(maximize), float(first_step), grad_scale))
    _call("b200_fused_sgd", _ptr(tab.recs), _ptr(tab.chunks), _i(tab.num_chunks), _ptr(tab.hyper), _stream())


def fused_adam(tables: dict, gi: int, ps, gs, ms, vs, steps, *, lr, beta1, beta2, eps, weight_decay, decoupled,
               maximize, grad_scale: float = 1.0) -> None:
    tab = _table(tables, gi, ps, gs, ms, vs)
    if tab.step_ptrs is None:
        ptrs = [s.data_ptr() if s.is_cuda else 0 for s in steps]
        tab.step_ptrs = torch.tensor(ptrs, dtype=torch.int64).to(ps[0].device)
        tab.hyper[8] = float(steps[0]) if steps else 0.0
    tab.set_hyper((lr, beta1, beta2, eps, weight_decay, float(decoupled), float(maximize), grad_scale))
    _call("b200_fused_adam", _ptr(tab.recs), _ptr(tab.chunks), _i(tab.num_chunks), _ptr(tab.hyper),
          _ptr(tab.step_ptrs), _i(tab.num_tensors), _stream(), kernels=2)
