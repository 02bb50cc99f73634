"""Ranking metrics: AUC, MRR, nDCG@k.

Behavioural spec: reference evaluation_functions.py — dcg_score :5-10, ndcg_score
:13-16, mrr_score :19-23, compute_amn :25-31, evaluation_split :33-46.  The numpy
functions keep the reference's names and conventions (gains ``2^y − 1``, discounts
``log2(i+2)``, descending argsort of the scores).  ``roc_auc_score`` is implemented
here with average ranks for ties, so scikit-learn is not a dependency.

``batched_ranking_metrics`` evaluates all impressions of a validation pass at once
on the device (one sort per metric), which is what ``Trainer.validate`` uses.
"""
from __future__ import annotations

from typing import Optional, Dict, Tuple

import numpy as np
import torch


def roc_auc_score(y_true, y_score) -> float:
    y_true = np.asarray(y_true).astype(np.float64).ravel()
    y_score = np.asarray(y_score).astype(np.float64).ravel()
    n_pos = float((y_true > 0).sum())
    n_neg = float(len(y_true) - n_pos)
    if n_pos == 0 or n_neg == 0:
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    order = np.argsort(y_score, kind="mergesort")
    s = y_score[order]
    ranks = np.empty(len(s), dtype=np.float64)
    i = 0
    while i < len(s):                     # average rank over each run of ties
        j = i
        while j + 1 < len(s) and s[j + 1] == s[i]:
            j += 1
        ranks[order[i:j + 1]] = 0.5 * (i + j) + 1.0
        i = j + 1
    return float((ranks[y_true > 0].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))


def dcg_score(y_true, y_score, k: int = 10) -> float:
    # kind="stable": upstream's bare np.argsort()[::-1] orders ties by whatever the installed numpy's unstable sort does
    # (SIMD sorts differ by CPU); the stable variant is its documented small-array behaviour — higher index first.
    order = np.argsort(y_score, kind="stable")[::-1]
    y_true = np.take(y_true, order[:k])
    gains = 2 ** y_true - 1
    discounts = np.log2(np.arange(len(y_true)) + 2)
    return float(np.sum(gains / discounts))


def ndcg_score(y_true, y_score, k: int = 10) -> float:
    return dcg_score(y_true, y_score, k) / dcg_score(y_true, y_true, k)


def mrr_score(y_true, y_score) -> float:
    order = np.argsort(y_score, kind="stable")[::-1]
    y_true = np.take(y_true, order)
    rr = y_true / (np.arange(len(y_true)) + 1)
    return float(np.sum(rr) / np.sum(y_true))


def compute_amn(y_true, y_score) -> Tuple[float, float, float, float]:
    return (roc_auc_score(y_true, y_score), mrr_score(y_true, y_score),
            ndcg_score(y_true, y_score, 5), ndcg_score(y_true, y_score, 10))


def evaluation_split(news_vecs, user_vecs, samples, nid2index, progress: bool = False) -> np.ndarray:
    """Full-ranking evaluation over precomputed vectors; one ``compute_amn`` row per sample."""
    rows = []
    it = range(len(samples))
    if progress:
        from tqdm import tqdm
        it = tqdm(it)
    for i in it:
        _, poss, negs, _, _ = samples[i]
        poss = list(poss) if isinstance(poss, (list, tuple)) else [poss]
        y_true = [1] * len(poss) + [0] * len(negs)
        idx = [nid2index[n] for n in poss + list(negs)]
        y_score = np.sum(np.multiply(news_vecs[idx], user_vecs[i]), axis=1)
        try:
            rows.append(compute_amn(y_true, y_score))
        except ValueError as e:           # single-class impression: skipped, as upstream
            print(e)
    return np.array(rows)


@torch.no_grad()
def batched_ranking_metrics(scores: torch.Tensor, labels: torch.Tensor,
                            valid: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
    """``scores``/``labels`` are ``[S, C]``; returns per-impression ``auc, mrr, ndcg5, ndcg10`` ``[S]``.

    ``valid`` (bool ``[S, C]``, true entries first in every row) makes the rows ragged: impressions with different
    candidate counts padded to a common width (full-ranking evaluation, evaluation_functions.py:33-46)."""
    scores = scores.float()
    labels = labels.to(scores.dtype)
    S, C = scores.shape
    if valid is not None:
        v = valid.to(scores.dtype)
        labels = labels * v
        scores = scores.masked_fill(~valid.bool(), float("-inf"))      # padding ranks after every real candidate
    n_pos = labels.sum(1)
    n_neg = (C if valid is None else v.sum(1)) - n_pos
    # AUC with ties counted 1/2: pairwise comparison (C is small: 5 in training, ≤ ~300 in eval)
    diff = scores[:, :, None] - scores[:, None, :]
    win = (diff > 0).to(scores.dtype) + 0.5 * (diff == 0).to(scores.dtype)
    pair = labels[:, :, None] * (1 - labels[:, None, :])
    if valid is not None:
        pair = pair * v[:, None, :]                                      # padding is nobody's negative
    auc = (win * pair).sum((1, 2)) / (n_pos * n_neg).clamp_min(1)
    # rank order of the reference: np.argsort(y_score)[::-1] (evaluation_functions.py:6,20), i.e. among equal scores the
    # HIGHER column comes first. Validation puts the positive in column 0, so a model whose scores saturate to equal
    # values ranks the positive last (MRR 1/C), not first: no inflated metrics for degenerate models.
    order = (C - 1) - scores.flip(1).argsort(dim=1, descending=True, stable=True)
    sorted_lab = labels.gather(1, order)
    pos = torch.arange(C, device=scores.device, dtype=scores.dtype)
    mrr = (sorted_lab / (pos + 1)).sum(1) / n_pos.clamp_min(1)
    disc = torch.log2(pos + 2)
    ideal = labels.sort(dim=1, descending=True).values

    def ndcg(k):
        k = min(k, C)
        dcg = ((2 ** sorted_lab[:, :k] - 1) / disc[:k]).sum(1)
        best = ((2 ** ideal[:, :k] - 1) / disc[:k]).sum(1)
        return dcg / best.clamp_min(1e-12)

    return {"auc": auc, "mrr": mrr, "ndcg5": ndcg(5), "ndcg10": ndcg(10)}
