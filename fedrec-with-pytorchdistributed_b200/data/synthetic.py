"""Synthetic MIND-shaped shards with a planted signal.

There is no network and no dataset on the build or GPU boxes, and BASELINE.json
fixes every run to synthetic data.  The generator reproduces the *file layout* of
the reference's ``UserData/``: ``news_index int64[N, 2, T]``
(token ids + attention mask, row 0 all zero = ``<unk>``), ``nid2index`` with
``'<unk>' -> 0``, and sample tuples ``(k, pos_nid, neg_nids, his_nids, uid)``.

Planted signal: every news item belongs to one of ``topics`` topics and draws its
title tokens mostly from that topic's slice of the vocabulary; every user prefers
a couple of topics; clicked history and the positive come from preferred topics,
negatives mostly from the rest.  A model that learns anything moves AUC/MRR/nDCG
away from chance, which is what the end-to-end tests check.
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass
from typing import Dict, List

import numpy as np


@dataclass
class SyntheticMind:
    news_index: np.ndarray
    nid2index: Dict[str, int]
    train_sam: List[tuple]
    valid_sam: List[tuple]
    news_topic: np.ndarray

    def save(self, data_dir: str) -> None:
        os.makedirs(data_dir, exist_ok=True)
        np.save(os.path.join(data_dir, "bert_news_index.npy"), self.news_index)
        for name, obj in (("bert_nid2index.pkl", self.nid2index), ("train_sam_uid.pkl", self.train_sam),
                          ("valid_sam_uid.pkl", self.valid_sam)):
            with open(os.path.join(data_dir, name), "wb") as f:
                pickle.dump(obj, f)


def make_synthetic_mind(num_news: int = 4096, num_train: int = 512, num_valid: int = 64,
                        title_len: int = 30, max_his_len: int = 50, topics: int = 16,
                        vocab_size: int = 30522, seed: int = 0, client: int = 0,
                        negs_per_impression: int = 20, zipf_a: float = 1.1,
                        min_his: int = 5, generic_frac: float = 0.25, liked_topics: int = 2) -> SyntheticMind:
    rng = np.random.default_rng(seed)                       # news table: shared by all clients
    urng = np.random.default_rng(seed * 1000003 + 17 * (client + 1))  # users: per client
    N, T = num_news, title_len
    topic = rng.integers(0, topics, size=N)
    topic[0] = 0
    # vocabulary: ids 1000.. are split into per-topic slices; 101/102 are [CLS]/[SEP]
    lo = 1000
    span = (vocab_size - lo) // topics
    lengths = rng.integers(max(4, T // 3), T - 1, size=N)
    tokens = np.zeros((N, T), dtype=np.int64)
    mask = np.zeros((N, T), dtype=np.int64)
    base = lo + topic * span
    body = base[:, None] + rng.integers(0, min(span, 512), size=(N, T))
    generic = rng.integers(lo, vocab_size, size=(N, T))
    use_generic = rng.random((N, T)) < generic_frac
    body = np.where(use_generic, generic, body)
    pos = np.arange(T)[None, :]
    inside = (pos >= 1) & (pos < lengths[:, None])
    tokens[:] = np.where(inside, body, 0)
    tokens[:, 0] = 101
    tokens[np.arange(N), lengths] = 102
    mask[:] = (pos <= lengths[:, None]).astype(np.int64)
    tokens[0] = 0
    mask[0] = 0
    news_index = np.stack([tokens, mask], axis=1)

    nids = ["<unk>"] + [f"N{100000 + i}" for i in range(1, N)]
    nid2index = {n: i for i, n in enumerate(nids)}

    # popularity: Zipf within topic so popular stories repeat inside a batch (dedup matters)
    by_topic = [np.flatnonzero((topic == t) & (np.arange(N) > 0)) for t in range(topics)]
    pop = [1.0 / np.power(np.arange(1, len(ix) + 1), zipf_a) for ix in by_topic]
    pop = [p / p.sum() for p in pop]

    def draw(tsel: np.ndarray, k: int) -> np.ndarray:
        out = np.empty(k, dtype=np.int64)
        ts = urng.choice(tsel, size=k)
        for j, t in enumerate(ts):
            out[j] = by_topic[t][urng.choice(len(by_topic[t]), p=pop[t])]
        return out

    def make(n: int, uid0: int) -> List[tuple]:
        sams = []
        for s in range(n):
            liked = urng.choice(topics, size=liked_topics, replace=False)
            other = np.setdiff1d(np.arange(topics), liked)
            L = int(urng.integers(min_his, max_his_len + max_his_len // 2))
            his = draw(liked, L)
            posn = int(draw(liked, 1)[0])
            n_neg = int(urng.integers(max(1, negs_per_impression // 4), negs_per_impression + 1))
            n_hard = n_neg // 8
            neg = np.concatenate([draw(other, n_neg - n_hard), draw(liked, n_hard)]) if n_hard else draw(other, n_neg)
            neg = neg[neg != posn]
            sams.append((s, nids[posn], [nids[i] for i in neg], [nids[i] for i in his],
                         f"U{client}_{uid0 + s}"))
        return sams

    return SyntheticMind(news_index, nid2index, make(num_train, 0), make(num_valid, num_train), topic)


def _main(argv=None):
    """``python -m fedrec_b200.data.synthetic --out UserData``: write a client shard in the upstream file layout."""
    import argparse
    ap = argparse.ArgumentParser(description=_main.__doc__)
    ap.add_argument("--out", default="UserData")
    ap.add_argument("--num_news", type=int, default=4096)
    ap.add_argument("--num_train", type=int, default=512)
    ap.add_argument("--num_valid", type=int, default=64)
    ap.add_argument("--title_len", type=int, default=30)
    ap.add_argument("--client", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args(argv)
    d = make_synthetic_mind(a.num_news, a.num_train, a.num_valid, a.title_len, seed=a.seed, client=a.client)
    d.save(a.out)
    print(f"wrote {a.out}: news_index {d.news_index.shape}, {len(d.train_sam)} train / {len(d.valid_sam)} valid impressions")


if __name__ == "__main__":
    _main()
