"""Sample GPU clocks / throttle reasons while a timed region runs (B200_PROFILING recipe).

Stand-alone on purpose: this file imports nothing from the package, so a process that must not
map any of this repository's native code (``bench.py --impl reference``) can load it by path.

Two back ends:

* NVML in-process (``pynvml`` from nvidia-ml-py): a thread polls SM clock, power and the clock
  event reasons every ``period_ms`` (2 ms by default), so even a 20 ms timed region holds several
  samples taken *inside* it;
* ``nvidia-smi -lms`` as a fallback where NVML cannot be loaded (coarser: tens of ms).
"""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional, Tuple

_FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
           "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
           "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

# NVML clock-event-reason bits (nvml.h: nvmlClocksEventReason*)
_REASON_BITS = {
    "sw_power_cap": 0x4,
    "hw_slowdown": 0x8,
    "sw_thermal_slowdown": 0x20,
    "hw_thermal_slowdown": 0x40,
}


def _physical_index(gpu_index: Optional[int]) -> Optional[int]:
    """CUDA ordinal -> NVML / nvidia-smi index when CUDA_VISIBLE_DEVICES is a plain index list."""
    if gpu_index is None:
        return None
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "").strip()
    if vis:
        parts = [p.strip() for p in vis.split(",") if p.strip()]
        if gpu_index < len(parts) and parts[gpu_index].isdigit():
            return int(parts[gpu_index])
    return gpu_index


class ClockSampler:
    def __init__(self, gpu_index: Optional[int] = None, period_ms: float = 2.0) -> None:
        self.gpu_index = _physical_index(gpu_index)
        self.period_ms = period_ms
        self.source = "none"
        self._proc: Optional[subprocess.Popen] = None
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        # (sm_mhz, sm_max_mhz, power_w, reasons bitmask)
        self._samples: List[Tuple[float, float, float, int]] = []

    # -- back ends -------------------------------------------------------------------------
    def _start_nvml(self) -> bool:
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu_index or 0)
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            return False
        try:
            get_reasons = nv.nvmlDeviceGetCurrentClocksEventReasons
        except AttributeError:
            get_reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons

        def poll():
            period = max(self.period_ms, 0.5) / 1e3
            while not self._stop.is_set():
                try:
                    sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1e3
                    rs = int(get_reasons(h))
                    self._samples.append((sm, mx, pw, rs))
                except Exception:
                    pass
                time.sleep(period)

        self._thread = threading.Thread(target=poll, daemon=True)
        self._thread.start()
        self.source = "nvml"
        return True

    def _start_smi(self) -> bool:
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return False
        cmd = [exe, f"--query-gpu={_FIELDS}", "--format=csv,noheader,nounits", "-lms", str(max(int(self.period_ms), 20))]
        if self.gpu_index is not None:
            cmd += ["-i", str(self.gpu_index)]
        try:
            self._proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return False

        def pump():
            assert self._proc is not None and self._proc.stdout is not None
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.strip().split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm, mx, pw = float(parts[1]), float(parts[2]), float(parts[3])
                except ValueError:
                    continue
                mask = 0
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                     parts[5:9]):
                    if val.lower().startswith("active"):
                        mask |= _REASON_BITS[name]
                self._samples.append((sm, mx, pw, mask))

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        self.source = "nvidia-smi"
        return True

    # -- public ----------------------------------------------------------------------------
    def start(self) -> "ClockSampler":
        if os.environ.get("B200DDP_CLOCKS", "nvml") != "smi" and self._start_nvml():
            return self
        self._start_smi()
        return self

    def mark(self) -> int:
        """Number of samples taken so far: pass it to ``stop`` / ``summary`` to summarise only what came after
        (start the sampler early and mark the beginning of the timed region)."""
        return len(self._samples)

    def wait_for_samples(self, n: int = 1, timeout_s: float = 3.0) -> None:
        t0 = time.time()
        while self.source != "none" and len(self._samples) < n and time.time() - t0 < timeout_s:
            time.sleep(0.005)

    def stop(self, first: int = 0, last: Optional[int] = None) -> Dict:
        self._stop.set()
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except Exception:
                self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self.summary(first, last)

    def summary(self, first: int = 0, last: Optional[int] = None) -> Dict:
        """Statistics over samples ``[first, last)``.  When that window is empty (region shorter than one
        sampling period) the three samples before it -- same load, warm-up -- are used instead and the
        result says so (``pre_region_fallback``)."""
        rows = self._samples[first:last]
        fallback = False
        if not rows:
            rows = self._samples[max(0, first - 3):first] or self._samples[-3:]
            fallback = True
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        sm = [r[0] for r in rows]
        power = [r[2] for r in rows]
        mask = 0
        for r in rows:
            mask |= r[3]
        reasons = sorted(n for n, b in _REASON_BITS.items() if mask & b)
        busy = [s for s, p in zip(sm, power) if p > 0.5 * max(power)] or sm
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(r[1] for r in rows), "reasons": reasons,
                "samples": len(rows), "power_w_max": max(power), "source": self.source,
                "pre_region_fallback": fallback}
