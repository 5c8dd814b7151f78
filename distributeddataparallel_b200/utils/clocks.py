"""Sample GPU clocks / throttle reasons while a timed region runs (B200_PROFILING recipe)."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
           "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
           "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: Optional[int] = None, period_ms: int = 100) -> None:
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        cmd = [exe, f"--query-gpu={_FIELDS}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms)]
        if self.gpu_index is not None:
            cmd += ["-i", str(self.gpu_index)]
        try:
            self._proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            assert self._proc is not None and self._proc.stdout is not None
            for line in self._proc.stdout:
                self._lines.append(line.strip())

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def mark(self) -> int:
        """Number of samples taken so far: pass it to ``stop`` / ``summary`` to summarise only what came after
        (start the sampler early -- nvidia-smi needs up to a second to come up on an 8-GPU box -- and mark the
        beginning of the timed region)."""
        return len(self._lines)

    def wait_for_samples(self, n: int = 1, timeout_s: float = 3.0) -> None:
        import time

        t0 = time.time()
        while self._proc is not None and len(self._lines) < n and time.time() - t0 < timeout_s:
            time.sleep(0.02)

    def stop(self, first: int = 0) -> Dict:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except Exception:
                self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self.summary(first)

    def summary(self, first: int = 0) -> Dict:
        sm, mx, power, reasons = [], [], [], set()
        lines = self._lines[first:]
        if not lines:                      # region shorter than one sampling period: the samples just before it
            lines = self._lines[-3:]       # (warm-up, same load) are the closest evidence there is
        for line in lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        busy = [s for s, p in zip(sm, power) if p > 0.5 * max(power)] or sm
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power)}
