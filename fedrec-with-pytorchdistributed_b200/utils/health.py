"""Straggler / failure detection and fault injection.

The reference's only knob is a two-day process-group timeout (client.py:227, server.py:59)
and its report states that one failed client fails the run. Here:

* ``Heartbeat`` publishes a per-rank progress counter to the job's key-value store and a
  watchdog thread on every rank flags peers whose counter has not moved for ``stale_s``
  seconds — reported as a clear error naming the rank instead of a silent multi-day hang;
* ``FaultInjector`` kills or delays a chosen rank at a chosen round (``FEDREC_FAULT``
  = ``kill:<rank>:<round>`` or ``delay:<rank>:<round>:<seconds>``) so the detector and the
  snapshot-resume path can be tested;
* resume itself is the snapshot mechanism of ``utils/checkpoint.py`` (round counter saved).
"""
from __future__ import annotations

import os
import threading
import time
from typing import Callable, Dict, List, Optional


class Heartbeat:
    def __init__(self, store, rank: int, world_size: int, period_s: float = 1.0, stale_s: float = 60.0,
                 on_stale: Optional[Callable[[List[int]], None]] = None, prefix: str = "fedrec/hb"):
        self.store, self.rank, self.W = store, rank, world_size
        self.period, self.stale_s = period_s, stale_s
        self.on_stale = on_stale
        self.prefix = prefix
        self.progress = 0
        self.stale: List[int] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._seen: Dict[int, tuple] = {}

    def beat(self, progress: Optional[int] = None) -> None:
        self.progress = self.progress + 1 if progress is None else progress
        self.store.set(f"{self.prefix}/{self.rank}", str(self.progress))

    def check_once(self, now: Optional[float] = None) -> List[int]:
        now = time.time() if now is None else now
        stale = []
        for r in range(self.W):
            if r == self.rank:
                continue
            try:
                val = self.store.get(f"{self.prefix}/{r}").decode() if self.store.check([f"{self.prefix}/{r}"]) else None
            except Exception:
                val = None
            last = self._seen.get(r)
            if last is None or last[0] != val:
                self._seen[r] = (val, now)
            elif now - last[1] > self.stale_s:
                stale.append(r)
        self.stale = stale
        return stale

    def _loop(self) -> None:
        while not self._stop.wait(self.period):
            stale = self.check_once()
            if stale and self.on_stale:
                self.on_stale(stale)

    def start(self) -> "Heartbeat":
        self.beat(0)
        self._thread = threading.Thread(target=self._loop, name="fedrec-heartbeat", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2 * self.period)


class FaultInjector:
    def __init__(self, spec: Optional[str] = None):
        spec = spec if spec is not None else os.environ.get("FEDREC_FAULT", "")
        self.kind, self.rank, self.round, self.arg = None, -1, -1, 0.0
        if spec:
            parts = spec.split(":")
            self.kind, self.rank, self.round = parts[0], int(parts[1]), int(parts[2])
            self.arg = float(parts[3]) if len(parts) > 3 else 0.0

    def maybe_fire(self, rank: int, rnd: int) -> None:
        if self.kind is None or rank != self.rank or rnd != self.round:
            return
        if int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") or 0) > 0:
            return          # the fault belongs to the first incarnation: after torchrun --max-restarts relaunched the job it runs clean
        if self.kind == "kill":
            print(f"[fedrec] fault injection: rank {rank} exits at round {rnd}", flush=True)
            os._exit(17)
        if self.kind == "delay":
            time.sleep(self.arg)
