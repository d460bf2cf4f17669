"""Device-side timing and clock sampling helpers (B200_PROFILING.md "Timing hygiene").

The reference only has wall-clock ``datetime.now()`` per iteration (example/main.py:77); every number this
framework reports is taken with CUDA events on the launching stream, bracketed by synchronisation, and
reduced with MAX over ranks.
"""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

import torch


class CudaTimer:
    """``with CudaTimer(stream) as t: ...`` -> ``t.ms`` (device time between the two events)."""

    def __init__(self, stream: Optional[torch.cuda.Stream] = None):
        self.stream = stream
        self.start = torch.cuda.Event(enable_timing=True)
        self.end = torch.cuda.Event(enable_timing=True)
        self.ms = 0.0

    def __enter__(self):
        torch.cuda.synchronize()
        self.start.record(self.stream) if self.stream is not None else self.start.record()
        return self

    def __exit__(self, *exc):
        self.end.record(self.stream) if self.stream is not None else self.end.record()
        self.end.synchronize()
        torch.cuda.synchronize()
        self.ms = self.start.elapsed_time(self.end)
        return False


_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks and throttle reasons in a background process while a timed region
    runs, and summarises them as ``{"sm_mhz": median under load, "sm_max_mhz": ..., "reasons": [...]}``."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "ClockSampler":
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index), "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return self

        def _pump():
            assert self.proc is not None and self.proc.stdout is not None
            for ln in self.proc.stdout:
                self.lines.append(ln.strip())

        self._thread = threading.Thread(target=_pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict:
        if self.proc is not None:
            time.sleep(self.period_ms / 1000.0 * 1.5)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        return self.summary()

    def summary(self) -> Dict:
        sm, sm_max, power = [], [], []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                sm_max.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        # "under load" = samples in the upper half of the observed power range
        lo, hi = min(power), max(power)
        loaded = [c for c, p in zip(sm, power) if p >= lo + 0.5 * (hi - lo)] or sm
        return {"sm_mhz": statistics.median(loaded), "sm_max_mhz": max(sm_max), "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": hi}
