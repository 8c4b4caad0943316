"""Device timing and clock sampling (the reference has no timers at all).

Every number is taken with CUDA events on the launching stream after a warm-up, bracketed by a
synchronize, and reduced as the max over ranks.  ``ClockSampler`` polls ``nvidia-smi`` during the
timed region so throttled runs can be rejected (B200_PROFILING.md).
"""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time
from typing import Callable, Dict, List, Optional

import torch


class CudaTimer:
    """``with CudaTimer() as t: ...`` -> ``t.ms`` (falls back to perf_counter on CPU)."""

    def __init__(self, stream: Optional["torch.cuda.Stream"] = None):
        self.cuda = torch.cuda.is_available()
        self.stream = stream
        self.ms = 0.0

    def __enter__(self):
        if self.cuda:
            torch.cuda.synchronize()
            self._a = torch.cuda.Event(enable_timing=True)
            self._b = torch.cuda.Event(enable_timing=True)
            self._a.record(self.stream)
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self._b.record(self.stream)
            torch.cuda.synchronize()
            self.ms = self._a.elapsed_time(self._b)
        else:
            self.ms = (time.perf_counter() - self._t0) * 1e3
        return False


_FLUSH = {}


def flush_l2(device=None, nbytes: int = 256 << 20):
    """Overwrite a buffer larger than the 126 MB L2 between timed iterations."""
    if not torch.cuda.is_available():
        return
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    buf = _FLUSH.get(dev)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _FLUSH[dev] = buf
    buf.zero_()


def time_kernel(fn: Callable[[], None], warmup: int = 3, iters: int = 10, flush: bool = True) -> Dict[str, float]:
    """Per-iteration CUDA-event timing with an L2 flush between iterations."""
    for _ in range(warmup):
        fn()
    times = []
    for _ in range(iters):
        if flush:
            flush_l2()
        with CudaTimer() as t:
            fn()
        times.append(t.ms)
    return {"min_ms": min(times), "median_ms": statistics.median(times), "max_ms": max(times), "iters": iters}


class ClockSampler:
    """Polls nvidia-smi every ``period_ms`` in a thread; ``summary()`` gives the JSON ``clocks`` block."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.rows: List[List[str]] = []
        self._proc = None
        self._thread = None

    def start(self):
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        try:
            self._proc = subprocess.Popen(
                [exe, f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-i", str(self.gpu_index), "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 9:
                    self.rows.append(parts)
        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self

    def summary(self) -> Dict:
        sm, smax, reasons, power = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2])); power.append(float(r[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power) if power else None}
