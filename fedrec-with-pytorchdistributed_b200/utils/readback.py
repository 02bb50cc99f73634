"""Lagged read-back of per-step scalars (loss, metrics) without stalling the launch queue.

``float(loss)`` right after a step makes the host wait for the whole step and then pay the launch latency of the next one
on the critical path (measured: ≈100 µs per 2.9 ms step). The reference does exactly that — it moves every gradient and
the loss to the CPU inside the batch loop (main.py:72-76, 90). ``LaggedScalars`` instead copies the value into a small
ring of pinned host slots on the step's own stream and hands it out ``lag`` pushes later, by which time the copy has long
finished; every value is still delivered, in order, and ``drain()`` returns whatever is still in flight.

On the CPU (or with ``lag=0``) it degenerates to a synchronous read, so callers need no special case.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, List, Optional, Tuple

import torch


class LaggedScalars:
    def __init__(self, device, lag: int = 1):
        self.device = torch.device(device)
        self.lag = max(int(lag), 0)
        self.cuda = self.device.type == "cuda" and self.lag > 0
        self._slots: Optional[torch.Tensor] = None
        self._free: Deque[int] = deque()
        self._inflight: Deque[Tuple[int, object]] = deque()          # (slot, event) on CUDA, (value, None) otherwise
        if self.cuda:
            self._slots = torch.zeros(self.lag + 1, dtype=torch.float32).pin_memory()
            self._free.extend(range(self.lag + 1))

    def push(self, value: torch.Tensor) -> Optional[float]:
        """Queue ``value`` (a 0-dim tensor); returns the value pushed ``lag`` calls ago, or None while the ring fills."""
        if not self.cuda:
            self._inflight.append((float(value), None))
        else:
            slot = self._free.popleft()
            self._slots[slot:slot + 1].copy_(value.detach().reshape(1).to(torch.float32), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self._inflight.append((slot, ev))
        return self._pop() if len(self._inflight) > self.lag else None

    def _pop(self) -> float:
        item, ev = self._inflight.popleft()
        if ev is None:
            return item
        ev.synchronize()
        out = float(self._slots[item])
        self._free.append(item)
        return out

    def drain(self) -> List[float]:
        """Everything still in flight, oldest first (call once after the loop)."""
        out = []
        while self._inflight:
            out.append(self._pop())
        return out
