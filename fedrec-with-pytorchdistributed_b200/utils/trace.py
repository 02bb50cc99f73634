"""Phase tracing: NVTX ranges + CUDA-event timers + Chrome trace export.

The reference has no profiler hooks at all (``UserModel.time`` is set and
never read). ``Tracer.span("name")`` brackets a phase with an NVTX range (visible in Nsight)
and a pair of CUDA events on the current stream; ``summary()`` aggregates device time per
phase after one synchronise, and ``export_chrome(path)`` writes a ``chrome://tracing`` file.
Disabled tracers cost one attribute check per span.
"""
from __future__ import annotations

import contextlib
import json
import time
from collections import defaultdict
from typing import Dict, List, Tuple

import torch


class Tracer:
    def __init__(self, enabled: bool = False, device=None):
        self.enabled = enabled
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self._events: List[Tuple[str, object, object, float]] = []

    @contextlib.contextmanager
    def span(self, name: str):
        if not self.enabled:
            yield
            return
        t0 = time.perf_counter()
        if self.cuda:
            torch.cuda.nvtx.range_push(name)
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
        try:
            yield
        finally:
            if self.cuda:
                e1.record()
                torch.cuda.nvtx.range_pop()
                self._events.append((name, e0, e1, t0))
            else:
                self._events.append((name, None, time.perf_counter() - t0, t0))

    def summary(self) -> Dict[str, Dict[str, float]]:
        if self.cuda and self._events:
            torch.cuda.synchronize()
        agg = defaultdict(lambda: {"count": 0, "total_ms": 0.0})
        for name, e0, e1, _ in self._events:
            ms = e0.elapsed_time(e1) if e0 is not None else e1 * 1e3
            agg[name]["count"] += 1
            agg[name]["total_ms"] += ms
        for v in agg.values():
            v["mean_ms"] = v["total_ms"] / max(v["count"], 1)
        return dict(agg)

    def export_chrome(self, path: str) -> None:
        if self.cuda and self._events:
            torch.cuda.synchronize()
        out = []
        for name, e0, e1, t0 in self._events:
            ms = e0.elapsed_time(e1) if e0 is not None else e1 * 1e3
            out.append({"name": name, "ph": "X", "ts": t0 * 1e6, "dur": ms * 1e3, "pid": 0, "tid": 0})
        with open(path, "w") as f:
            json.dump({"traceEvents": out}, f)

    def reset(self) -> None:
        self._events.clear()
