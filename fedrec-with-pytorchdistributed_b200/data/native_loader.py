"""The batch producer of ``csrc/host_loader.cu`` behind the ``BatchSampler`` interface.

A C++ thread draws the impressions (fresh permutation per pass), samples the negatives and packs the ids in the layout
the training step reads; ``next_packed`` hands out the next prepared batch. The launch loop then costs one 11 KB copy per
step instead of 130 us of numpy (``BatchSampler``) or upstream's per-item ``__getitem__`` + collate (dataset.py:69-86).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import numpy as np
import torch

from .dataset import ImpressionTable, npratio as _npratio


class NativeBatchLoader:
    def __init__(self, table: ImpressionTable, batch_size: int, npratio: int = _npratio, seed: int = 0, depth: int = 8):
        from ..ops.native import lib
        self._lib = lib()
        if self._lib is None:
            raise RuntimeError("libfedrec_b200.so is not available")
        if len(table) < 1:
            raise ValueError("empty shard")
        self.B, self.C, self.L = batch_size, 1 + npratio, table.max_his_len
        self.n = self.B * (self.L + self.C)
        pos = np.ascontiguousarray(table.pos, dtype=np.int32)
        his = np.ascontiguousarray(table.his, dtype=np.int32)
        ptr = np.ascontiguousarray(table.neg_ptr, dtype=np.int64)
        idx = np.ascontiguousarray(table.neg_idx, dtype=np.int32)
        as_p = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
        self._h = self._lib.fr_loader_create(as_p(pos), as_p(his), as_p(ptr), as_p(idx) if idx.size else None, len(table),
                                             self.L, self.B, npratio, int(seed) & (2 ** 64 - 1), depth)
        if not self._h:
            raise RuntimeError(self._lib.fr_last_error().decode())
        assert self._lib.fr_loader_batch_ints(ctypes.c_void_p(self._h)) == self.n
        self._buf = np.zeros(self.n, dtype=np.int32)
        self._epoch = ctypes.c_longlong(0)
        self.epoch = 0

    def next_packed(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        """The next batch as int32 ``[B*L + B*C]`` (histories first). Without ``out`` the same internal array is
        returned every call: consume it (``TrainStep.run_host`` copies it at once) before asking for the next."""
        dst = self._buf if out is None else out
        assert dst.dtype == np.int32 and dst.size >= self.n and dst.flags["C_CONTIGUOUS"]
        rc = self._lib.fr_loader_next(ctypes.c_void_p(self._h), dst.ctypes.data_as(ctypes.c_void_p),
                                      ctypes.byref(self._epoch))
        if rc != 0:
            raise RuntimeError(self._lib.fr_last_error().decode())
        self.epoch = int(self._epoch.value)
        return dst

    def next(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """``(cand [B, C], his [B, L])`` int32 CPU tensors, like ``BatchSampler.next`` (fresh storage)."""
        ids = torch.from_numpy(self.next_packed().copy())
        return ids[self.B * self.L:].view(self.B, self.C), ids[: self.B * self.L].view(self.B, self.L)

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fr_loader_destroy(ctypes.c_void_p(self._h))
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
