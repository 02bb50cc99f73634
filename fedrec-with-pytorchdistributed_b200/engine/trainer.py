"""Reference-compatible training-loop API.

Reference: ``process_news_grad`` main.py:20-42, ``process_user_grad`` main.py:48-52,
``train_on_step`` main.py:55-91 (client variant client.py:61-101), ``Trainer``
main.py:95-139 / client.py:105-189 (``validate`` :149-171). The four upstream copies of
this code are one module here; the names and call signatures are kept so scripts written
against the reference run, but each call lands on the device-side engine:

* ``process_news_grad`` is a device segment-sum keyed by news row (``ops.segment_sum``);
* ``train_on_step`` runs one fused step per batch (gradient → aggregate → Adam) instead of
  accumulating news gradients over an epoch and stepping once;
* ``Trainer`` wraps no DDP: gradient exchange is the aggregator's job; snapshots are
  rank-0, atomic, and store the *next* epoch;
* ``validate`` reports means over all impressions, computed on device.
"""
from __future__ import annotations

import os
from typing import Dict, Iterable, Optional, Tuple

import numpy as np
import torch

from .. import ops
from ..utils import checkpoint as ckpt
from ..utils.metrics import batched_ranking_metrics
from .step import TrainStep


def process_news_grad(candidate_info, his_info) -> Dict[int, torch.Tensor]:
    """``{news row: summed grad[D]}`` from ``[ids, vecs, grads]`` triples (vecs are ignored, as upstream)."""
    cand_ids, _, cand_grad = candidate_info
    his_ids, _, his_grad = his_info
    D = cand_grad.shape[-1]
    ids = torch.cat([torch.as_tensor(np.asarray(his_ids)).reshape(-1), torch.as_tensor(np.asarray(cand_ids)).reshape(-1)])
    grads = torch.cat([torch.as_tensor(his_grad).reshape(-1, D), torch.as_tensor(cand_grad).reshape(-1, D)])
    ids = ids.to(grads.device)
    uniq, inv = torch.unique(ids, return_inverse=True)
    summed = ops.segment_sum(grads.float(), inv, uniq.numel())
    return {int(i): g for i, g in zip(uniq.tolist(), summed)}


def process_user_grad(model_param: Iterable[Tuple[str, torch.nn.Parameter]]) -> Dict[str, torch.Tensor]:
    return {name: p.grad for name, p in model_param}


def train_on_step(model, train_dl, optimizer=None, dp_enabled: bool = False, sigma: Optional[float] = None,
                  aggregator=None, step: Optional[TrainStep] = None):
    """One pass over ``train_dl``; every batch is a full step. Returns the summed loss (main.py:91)."""
    model = getattr(model, "module", model)
    step = step or getattr(model, "_train_step", None) or TrainStep(model)
    model._train_step = step
    old_noise = step.cfg.news_grad_noise
    if dp_enabled and sigma:
        step.cfg.news_grad_noise = float(sigma)
    total = torch.zeros((), device=model.device)
    try:
        for batch in train_dl:
            cand, his = batch[0], batch[1]
            cand = torch.as_tensor(np.asarray(cand) if not isinstance(cand, torch.Tensor) else cand).to(model.device)
            his = torch.as_tensor(np.asarray(his) if not isinstance(his, torch.Tensor) else his).to(model.device)
            loss = step.run(cand, his, train=True)
            if aggregator is not None:
                aggregator.allreduce_mean_()
            model.adam_step()
            model.cnt += 1
            total += loss
    finally:
        step.cfg.news_grad_noise = old_noise
    return total


@torch.no_grad()
def validate_impressions(step: TrainStep, model, valid_sam, nid2index, batch: int = 256, last_only: bool = False):
    """Mean loss / AUC / MRR / nDCG@5 / nDCG@10 over validation impressions, computed on the device.

    Candidates per impression: the positive + the *last* ``npratio`` negatives (client.py:160). Upstream returns the last
    sample's metrics only, rounded to two decimals, next to the mean loss (client.py:171); means are reported here
     unless ``last_only=True`` asks for the upstream numbers.
    """
    cfg = model.cfg
    L, K = cfg.max_his_len, cfg.npratio
    cands, hiss = [], []
    for s in valid_sam:
        _, pos, neg, his, _ = s
        pos = pos[0] if isinstance(pos, (list, tuple)) else pos
        negs = list(neg)[-K:]
        negs = negs + ["<unk>"] * (K - len(negs))
        cands.append([nid2index[pos]] + [nid2index[n] for n in negs])
        h = [nid2index[n] for n in his][-L:]
        hiss.append(h + [0] * (L - len(h)))
    if not cands:
        return 0.0, 0.0, 0.0, 0.0, 0.0
    cand = torch.tensor(cands, device=model.device)
    his = torch.tensor(hiss, device=model.device)
    scores = torch.cat([step.scores(cand[i:i + batch], his[i:i + batch]) for i in range(0, len(cands), batch)])
    labels = torch.zeros_like(scores)
    labels[:, 0] = 1
    loss = torch.nn.functional.cross_entropy(scores, torch.zeros(len(cands), dtype=torch.long, device=scores.device))
    m = batched_ranking_metrics(scores, labels)
    if last_only:
        return (float(loss),) + tuple(round(float(m[k][-1]), 2) for k in ("auc", "mrr", "ndcg5", "ndcg10"))
    return (float(loss), float(m["auc"].mean()), float(m["mrr"].mean()), float(m["ndcg5"].mean()),
            float(m["ndcg10"].mean()))


class Trainer:
    def __init__(self, model, train_data, optimizer=None, save_every: int = 1, snapshot_path: str = "snapshot.pt",
                 isTrain: bool = True, aggregator=None, logger=None):
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.global_rank = int(os.environ.get("RANK", "0"))
        self.model = getattr(model, "module", model)
        self.train_data = train_data
        self.optimizer = optimizer            # accepted and ignored, as upstream (zero_grad-only SGD)
        self.save_every = max(int(save_every), 1)
        self.snapshot_path = snapshot_path
        self.epochs_run = 0
        self.aggregator = aggregator if isTrain else None
        self.logger = logger
        self.step = TrainStep(self.model)
        self.model._train_step = self.step
        if snapshot_path and os.path.exists(snapshot_path):
            self._load_snapshot(snapshot_path)

    def _load_snapshot(self, path: str) -> None:
        info = ckpt.load_snapshot(path, self.model, map_location="cpu")
        self.epochs_run = info["EPOCHS_RUN"]
        print(f"Resuming training from snapshot at Epoch {self.epochs_run}")

    def _save_snapshot(self, epoch: int) -> None:
        if ckpt.save_snapshot(self.snapshot_path, self.model, next_epoch=epoch + 1, rank=self.global_rank):
            print(f"Epoch {epoch} | Training snapshot saved at {self.snapshot_path}")

    def validate(self, valid_sam, nid2index, news_index=None, device=None, batch: int = 256, last_only: bool = False):
        """Mean loss / AUC / MRR / nDCG@5 / nDCG@10 over validation impressions (see ``validate_impressions``)."""
        return validate_impressions(self.step, self.model, valid_sam, nid2index, batch, last_only)

    def train(self, max_epochs: int, valid_sam=None, nid2index=None, news_index=None, device=None,
              dp_enabled: bool = False, sigma: Optional[float] = None):
        last = None
        for epoch in range(self.epochs_run, max_epochs):
            loss = train_on_step(self.model, self.train_data, self.optimizer, dp_enabled, sigma,
                                 aggregator=self.aggregator, step=self.step)
            last = loss
            rec = {"epoch": epoch, "training_loss": float(loss)}
            if valid_sam is not None and nid2index is not None:
                v = self.validate(valid_sam, nid2index, news_index, device)
                rec.update({"validation_loss": v[0], "valid_auc": v[1], "valid_mrr": v[2],
                            "val_ndcg@5": v[3], "val_ndcg@10": v[4]})
            if self.logger is not None:
                self.logger.log(rec, step=epoch)
            if self.snapshot_path and (epoch + 1) % self.save_every == 0:
                self._save_snapshot(epoch)
            self.epochs_run = epoch + 1
        return last
