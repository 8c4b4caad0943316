"""Opt-in tracing (the reference has no timers or listeners).

``DAPB_TRACE=1`` turns :func:`span` into NVTX ranges (visible in Nsight tools) and records wall / CUDA-event
durations per span name; ``summary()`` returns them.  Disabled, a span costs one dict lookup.
"""
from __future__ import annotations

import os
import time
from collections import defaultdict
from contextlib import contextmanager
from typing import Dict

ENABLED = os.environ.get("DAPB_TRACE", "0") not in ("", "0")
_STATS: Dict[str, list] = defaultdict(lambda: [0, 0.0])


@contextmanager
def span(name: str, sync: bool = False):
    """``with span("kmeans.step"): ...`` -- NVTX range + accumulated host time (device time if ``sync``)."""
    if not ENABLED:
        yield
        return
    import torch
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.nvtx.range_push(name)
        if sync:
            torch.cuda.synchronize()
    t0 = time.perf_counter()
    try:
        yield
    finally:
        if cuda and sync:
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        st = _STATS[name]
        st[0] += 1
        st[1] += dt
        if cuda:
            torch.cuda.nvtx.range_pop()


def summary() -> Dict[str, Dict[str, float]]:
    return {k: {"calls": v[0], "total_s": v[1], "mean_ms": 1e3 * v[1] / max(v[0], 1)} for k, v in _STATS.items()}


def reset():
    _STATS.clear()
