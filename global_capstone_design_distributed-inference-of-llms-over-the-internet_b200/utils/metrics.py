"""Structured metrics: JSONL step/hop records, latency histograms, NVTX ranges, CUDA-event timers.

The reference only keeps ``time.perf_counter()`` lists on the client (src/rpc_transport.py:98-103,
760-769; src/main.py:213-225) and logs free-form INFO lines.  Here every component can emit
machine-readable records (``PETALS_B200_METRICS=/path/file.jsonl``), summarise latencies as
p50/p90/p99, wrap regions in NVTX ranges for nsys/ncu timelines, and time device work with CUDA events
(the only clock that is valid for GPU numbers).
"""
from __future__ import annotations

import contextlib
import json
import os
import threading
import time
from typing import Dict, Iterable, List, Optional

_lock = threading.Lock()
_sink = None


def _open_sink():
    global _sink
    path = os.environ.get("PETALS_B200_METRICS")
    if path and _sink is None:
        _sink = open(path, "a", buffering=1)
    return _sink


def emit(kind: str, **fields) -> None:
    """Append one JSON line ``{"ts": ..., "kind": kind, ...}`` when a metrics file is configured."""
    f = _open_sink()
    if f is None:
        return
    rec = {"ts": round(time.time(), 6), "kind": kind, **fields}
    with _lock:
        f.write(json.dumps(rec, default=str) + "\n")


def percentiles(samples: Iterable[float], qs=(50, 90, 99)) -> Dict[str, float]:
    s = sorted(samples)
    if not s:
        return {}
    out = {"n": len(s), "mean": sum(s) / len(s), "min": s[0], "max": s[-1]}
    for q in qs:
        out[f"p{q}"] = s[min(len(s) - 1, int(round(q / 100.0 * (len(s) - 1))))]
    return out


class LatencyHistogram:
    """Log-spaced latency histogram (microseconds .. seconds) with cheap inserts."""

    EDGES_MS = [0.05, 0.1, 0.2, 0.5, 1, 2, 5, 10, 20, 50, 100, 200, 500, 1000, 5000]

    def __init__(self):
        self.counts = [0] * (len(self.EDGES_MS) + 1)
        self.samples: List[float] = []

    def add(self, seconds: float) -> None:
        ms = seconds * 1e3
        i = 0
        while i < len(self.EDGES_MS) and ms > self.EDGES_MS[i]:
            i += 1
        self.counts[i] += 1
        if len(self.samples) < 100000:
            self.samples.append(ms)

    def summary(self) -> Dict[str, float]:
        return percentiles(self.samples)


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range when CUDA is present (shows up in nsys / ncu timelines); no-op otherwise."""
    try:
        import torch

        on = torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        on = False
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class CudaTimer:
    """Device-time a region with CUDA events on the current stream: ``with CudaTimer() as t: ...; t.ms``."""

    def __init__(self, sync: bool = True):
        import torch

        self._torch = torch
        self.sync = sync
        self.ms: Optional[float] = None

    def __enter__(self):
        t = self._torch
        self._e0, self._e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
        self._e0.record()
        return self

    def __exit__(self, *exc):
        self._e1.record()
        if self.sync:
            self._e1.synchronize()
            self.ms = self._e0.elapsed_time(self._e1)
        return False
