"""Tracing helpers: NVTX ranges (visible in Nsight tools) and per-batch timestamps into the JSONL metrics sink.

The reference's only instrumentation is one wall-clock line per job (worker.py:105,219).  Here every
batch can be bracketed with `span("name")`, which pushes an NVTX range when CUDA is available and
records host timestamps into a `Metrics` object (utils/metrics.py) for p50/p99 summaries.
"""
from __future__ import annotations

import contextlib
import time

try:
    import torch
    _NVTX = torch.cuda.nvtx if torch.cuda.is_available() else None
except Exception:  # pragma: no cover
    _NVTX = None


@contextlib.contextmanager
def span(name: str, metrics=None, **fields):
    t0 = time.perf_counter()
    if _NVTX is not None:
        _NVTX.range_push(name)
    try:
        yield
    finally:
        if _NVTX is not None:
            _NVTX.range_pop()
        dt = (time.perf_counter() - t0) * 1e3
        if metrics is not None:
            metrics.observe(name + "_ms", dt)
            metrics.emit(name, ms=round(dt, 4), **fields)


class DeviceTimer:
    """CUDA-event stopwatch on the current stream (device time, not wall clock)."""

    def __init__(self):
        import torch as _t
        self._t = _t
        self.a = _t.cuda.Event(enable_timing=True)
        self.b = _t.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.a.record()
        return self

    def __exit__(self, *exc):
        self.b.record()

    def ms(self) -> float:
        self.b.synchronize()
        return self.a.elapsed_time(self.b)
