"""Impression sampling and news lookup datasets.

Behavioural spec: reference dataset.py (TrainDataset :69-86, newsample :10-14,
NewsDataset :22-30, NewsPartDataset :36-46, NewsUpdatorDataset :52-63,
UserDataset :92-109).  Differences adopted on purpose:
histories are truncated to the most recent ``max_his_len`` clicks, sampling is
driven by a seedable generator, and a client owns its whole shard.

Two layers live here:

* the map-style ``Dataset`` classes with the reference's names and item layout,
  for users who bring their own ``DataLoader``;
* ``ImpressionTable`` + ``BatchSampler``: the shard pre-indexed into flat int32
  arrays so a whole round's batch is produced by vectorised numpy into a pinned
  staging buffer, ready for one async H2D copy.
"""
from __future__ import annotations

import os
import pickle
import random
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset

npratio = 4
max_his_len = 50
UNK = "<unk>"


def newsample(negatives: Sequence, ratio: int, rng: Optional[random.Random] = None) -> List:
    """Pick ``ratio`` negatives; pad with ``<unk>`` when there are too few."""
    negatives = list(negatives)
    if ratio > len(negatives):
        return negatives + [UNK] * (ratio - len(negatives))
    return (rng or random).sample(negatives, ratio)


class NewsDataset(Dataset):
    """Row ``i`` is the ``[2, T]`` (token ids, attention mask) record of news ``i``."""

    def __init__(self, news_index):
        self.news_index = news_index

    def __len__(self):
        return len(self.news_index)

    def __getitem__(self, idx):
        return self.news_index[idx]


class NewsPartDataset(Dataset):
    """``(nid, news_index[nid])`` for a list of news ids."""

    def __init__(self, news_index, nids):
        self.news_index = news_index
        self.nids = nids

    def __len__(self):
        return len(self.nids)

    def __getitem__(self, idx):
        nid = self.nids[idx]
        return nid, self.news_index[nid]


class NewsUpdatorDataset(Dataset):
    """``(news_index[nid], grad)`` pairs feeding the text-head update."""

    def __init__(self, news_index, news_ids, news_grads):
        assert len(news_ids) == len(news_grads)
        self.news_index = news_index
        self.news_ids = news_ids
        self.news_grads = news_grads

    def __len__(self):
        return len(self.news_ids)

    def __getitem__(self, idx):
        return self.news_index[self.news_ids[idx]], self.news_grads[idx]


def _pad_history(his_idx: Sequence[int], length: int) -> List[int]:
    his_idx = list(his_idx)[-length:]
    return his_idx + [0] * (length - len(his_idx))


class TrainDataset(Dataset):
    """One impression → ``(candidates int64[1+npratio], history int64[L], label 0)``.

    The positive is always candidate 0, hence the constant label.
    """

    def __init__(self, train_sam, nid2index: Dict, news_index, npratio: int = npratio,
                 max_his_len: int = max_his_len, seed: Optional[int] = None):
        self.news_index = news_index
        self.nid2index = nid2index
        self.train_sam = train_sam
        self.npratio = npratio
        self.max_his_len = max_his_len
        self._rng = random.Random(seed) if seed is not None else None

    def __len__(self):
        return len(self.train_sam)

    def __getitem__(self, idx):
        _, pos, neg, his, _ = self.train_sam[idx]
        neg = newsample(neg, self.npratio, self._rng)
        cand = np.array([self.nid2index[n] for n in [pos] + neg], dtype=np.int64)
        his = np.array(_pad_history([self.nid2index[n] for n in his], self.max_his_len), dtype=np.int64)
        return cand, his, np.array(0, dtype=np.int64)


class UserDataset(Dataset):
    """History lookup into *precomputed* news vectors (inference path, dataset.py:92-109)."""

    def __init__(self, args, samples, news_vecs, nid2index):
        self.samples = samples
        self.max_his_len = getattr(args, "max_his_len", max_his_len) if args is not None else max_his_len
        self.news_vecs = news_vecs
        self.nid2index = nid2index

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        _, _, _, his, _ = self.samples[idx]
        his = _pad_history([self.nid2index[n] for n in his], self.max_his_len)
        return self.news_vecs[his]


# ---------------------------------------------------------------------------------
# Vectorised path
# ---------------------------------------------------------------------------------
class ImpressionTable:
    """A client's shard as flat arrays.

    ``pos[S]``, ``his[S, L]`` (0-padded, most recent L), and the negatives as a CSR
    pair ``neg_ptr[S+1]`` / ``neg_idx[*]``.  All int32 news-table row numbers.
    """

    def __init__(self, samples, nid2index: Dict, max_his_len: int = max_his_len):
        S = len(samples)
        self.max_his_len = max_his_len
        self.pos = np.zeros(S, dtype=np.int32)
        self.his = np.zeros((S, max_his_len), dtype=np.int32)
        self.his_len = np.zeros(S, dtype=np.int32)
        ptr = [0]
        flat: List[int] = []
        for s, sam in enumerate(samples):
            _, pos, neg, his, _ = sam
            poss = pos if isinstance(pos, (list, tuple)) else [pos]
            self.pos[s] = nid2index[poss[0]]
            h = [nid2index[n] for n in his][-max_his_len:]
            self.his[s, : len(h)] = h
            self.his_len[s] = len(h)
            flat.extend(nid2index[n] for n in neg)
            ptr.append(len(flat))
        self.neg_ptr = np.asarray(ptr, dtype=np.int64)
        self.neg_idx = np.asarray(flat, dtype=np.int32)

    def __len__(self):
        return len(self.pos)


class BatchSampler:
    """Draws rounds of ``B`` impressions into pinned int32 staging buffers.

    ``next()`` returns ``(cand[B, C], his[B, L])`` views of the *same* pinned
    buffers every call (double-buffered), so the caller can issue
    ``device.copy_(non_blocking=True)`` without allocating.
    """

    def __init__(self, table: ImpressionTable, batch_size: int, npratio: int = npratio,
                 seed: int = 0, pin: Optional[bool] = None, drop_last: bool = True):
        self.table = table
        self.B = batch_size
        self.C = 1 + npratio
        self.npratio = npratio
        self.rng = np.random.default_rng(seed)
        pin = torch.cuda.is_available() if pin is None else pin
        L = table.max_his_len
        self._bufs = []
        for _ in range(2):
            cand = torch.zeros(batch_size, self.C, dtype=torch.int32)
            his = torch.zeros(batch_size, L, dtype=torch.int32)
            if pin:
                cand, his = cand.pin_memory(), his.pin_memory()
            self._bufs.append((cand, his))
        self._flip = 0
        self._last = 0
        self._events = [None, None]          # CUDA event of the last async copy out of each staging buffer
        self._order = np.zeros(0, dtype=np.int64)
        self._cursor = 0
        self.epoch = 0

    def _next_indices(self) -> np.ndarray:
        S = len(self.table)
        out = []
        need = self.B
        while need > 0:
            if self._cursor >= len(self._order):
                self._order = self.rng.permutation(S)
                self._cursor = 0
                self.epoch += 1
            take = self._order[self._cursor: self._cursor + need]
            self._cursor += len(take)
            need -= len(take)
            out.append(take)
        return np.concatenate(out)

    def next(self) -> Tuple[torch.Tensor, torch.Tensor]:
        idx = self._next_indices()
        ev = self._events[self._flip]
        if ev is not None:                   # the host must not overwrite a pinned buffer an H2D copy still reads
            ev.synchronize()
            self._events[self._flip] = None
        cand, his = self._bufs[self._flip]
        self._last = self._flip
        self._flip ^= 1
        t = self.table
        cand_np = cand.numpy()
        cand_np[:, 0] = t.pos[idx]
        lo = t.neg_ptr[idx]
        cnt = t.neg_ptr[idx + 1] - lo
        # distinct negatives per row when there are enough, else all of them + <unk>(0) padding. Floyd's sampling run on
        # all rows at once: draw k picks t from [0, n-K+k]; a value already taken is replaced by n-K+k, which no
        # earlier draw can have produced (K numpy passes per batch instead of one generator call per impression)
        K = self.npratio
        enough = cnt >= K
        picks = np.zeros((self.B, K), dtype=np.int64)
        u = self.rng.random((self.B, K))
        for k in range(K):
            top = np.maximum(cnt - K + k, 0)
            tk = np.minimum((u[:, k] * (top + 1)).astype(np.int64), top)
            if k:
                tk = np.where((picks[:, :k] == tk[:, None]).any(axis=1), top, tk)
            picks[:, k] = tk
        if K:
            order = self.rng.random((self.B, K)).argsort(axis=1)       # Floyd's output order is biased; shuffle the picks
            picks = np.take_along_axis(picks, order, axis=1)
        few = np.arange(K)[None, :] < cnt[:, None]
        off = np.where(enough[:, None], picks, np.arange(K)[None, :])
        valid = enough[:, None] | few
        src = np.minimum(lo[:, None] + off, max(len(t.neg_idx) - 1, 0))
        vals = t.neg_idx[src] if len(t.neg_idx) else np.zeros_like(src, dtype=np.int32)
        cand_np[:, 1:] = np.where(valid, vals, 0)
        his.numpy()[:] = t.his[idx]
        return cand, his

    def copy_to(self, cand_dev: torch.Tensor, his_dev: torch.Tensor) -> None:
        """Async H2D copy of the batch returned by the last ``next()``; remembers when the staging buffer is free again."""
        cand, his = self._bufs[self._last]
        cand_dev.copy_(cand, non_blocking=True)
        his_dev.copy_(his, non_blocking=True)
        if cand_dev.is_cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(cand_dev.device))
            self._events[self._last] = ev

    def __iter__(self):
        while True:
            yield self.next()


class DevicePrefetcher:
    """Keeps the next round's ids in flight: the H2D copy of batch i+1 runs on its own stream while step i computes.

    ``next()`` returns device tensors ``(cand[B, C], his[B, L])`` that stay valid until the following ``next()``;
    two device slots and two pinned slots alternate, ordered by CUDA events (no host synchronisation in steady state).
    """

    def __init__(self, sampler: BatchSampler, device: torch.device):
        self.sampler, self.device = sampler, torch.device(device)
        self.cuda = self.device.type == "cuda"
        L = sampler.table.max_his_len
        self._slots = [(torch.zeros(sampler.B, sampler.C, dtype=torch.int32, device=self.device),
                        torch.zeros(sampler.B, L, dtype=torch.int32, device=self.device)) for _ in range(2)]
        self._stream = torch.cuda.Stream(device=self.device) if self.cuda else None
        self._ready = [None, None]           # copy finished (consumer waits)
        self._free = [None, None]            # consumer finished with the slot (copy stream waits)
        self._i = 0
        self._issue(0)

    def _issue(self, slot: int) -> None:
        self.sampler.next()
        c, h = self._slots[slot]
        if not self.cuda:
            self.sampler.copy_to(c, h)
            return
        with torch.cuda.stream(self._stream):
            if self._free[slot] is not None:
                self._stream.wait_event(self._free[slot])
            self.sampler.copy_to(c, h)
            ev = torch.cuda.Event()
            ev.record(self._stream)
            self._ready[slot] = ev

    def next(self):
        slot = self._i & 1
        if self.cuda:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self._ready[slot])
            prev = slot ^ 1                  # the slot handed out last time is free once work queued so far is done
            ev = torch.cuda.Event()
            ev.record(cur)
            self._free[prev] = ev
        self._i += 1
        out = self._slots[slot]
        self._issue(slot ^ 1)
        return out


def _floyd_sample(rng, n: int, k: int) -> np.ndarray:
    """k distinct integers from range(n) without materialising a permutation."""
    chosen: List[int] = []
    seen = set()
    for j in range(n - k, n):
        t = int(rng.integers(0, j + 1))
        if t in seen:
            t = j
        seen.add(t)
        chosen.append(t)
    return np.asarray(chosen, dtype=np.int64)


def load_user_data(data_dir: str):
    """Read the four files of a client shard (layout: SURVEY.md C29, main.py:146-157)."""
    with open(os.path.join(data_dir, "bert_nid2index.pkl"), "rb") as f:
        nid2index = pickle.load(f)
    news_index = np.load(os.path.join(data_dir, "bert_news_index.npy"), allow_pickle=True)
    with open(os.path.join(data_dir, "train_sam_uid.pkl"), "rb") as f:
        train_sam = pickle.load(f)
    with open(os.path.join(data_dir, "valid_sam_uid.pkl"), "rb") as f:
        valid_sam = pickle.load(f)
    return nid2index, news_index, train_sam, valid_sam
