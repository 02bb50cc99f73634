"""Serving: top-k news for a batch of click histories from precomputed news vectors.

Upstream stops at validation over sampled candidate lists (main.py:93-132) and only sketches the precomputed-vector
path (``UserDataset``, dataset.py:92-109; ``evaluation_split``, evaluation_functions.py:33-46). A deployed recommender
does the other half of that sketch: the news table is encoded once per model version (frozen trunk + head, on the sm_100a
kernels when a GPU is present) and kept resident — 65,536 news x 400 floats = 105 MB of the 180 GB — and a request is the
user encoder over ``[B, L]`` gathered vectors plus one ``[B, D] x [D, N]`` product and a top-k. The scores are the
training-time ones (``sigmoid(news . user)``, model.py:118-119), so ranking a candidate list here equals ranking it with
``UserModel.forward``.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence, Tuple

import torch

from ..models.model import UserModel, _as_ids


class Recommender:
    def __init__(self, model: UserModel, table_dtype: Optional[torch.dtype] = None, encode_batch: int = 2048):
        self.model = model
        self.device = model.device
        self.L = model.cfg.max_his_len
        self.encode_batch = encode_batch
        self.table_dtype = table_dtype or torch.float32
        self.news_vecs: Optional[torch.Tensor] = None      # [N, D] fp32 (user-encoder input)
        self._table: Optional[torch.Tensor] = None          # [N, D] in table_dtype (scoring operand)
        self.version = 0

    # ---- table ----------------------------------------------------------------------------------------------
    @torch.no_grad()
    def refresh(self, news_vecs: Optional[torch.Tensor] = None) -> "Recommender":
        """(Re)encode the news table after the model changed (a new snapshot, a finished federated round), or install
        vectors computed elsewhere."""
        if hasattr(self.model, "sync_optimizer"):
            self.model.sync_optimizer()
        v = self.model.encode_all_news(self.encode_batch) if news_vecs is None else news_vecs.to(self.device, torch.float32)
        if v.shape[0] != self.model.news_tokens.shape[0]:
            raise ValueError(f"news table has {self.model.news_tokens.shape[0]} rows, got vectors for {v.shape[0]}")
        self.news_vecs = v
        self._table = v.to(self.table_dtype)
        self.version += 1
        return self

    def save_table(self, path: str) -> None:
        """The encoded table as a file, so a serving process starts without running the text encoder."""
        self._need_table()
        tmp = path + ".tmp"
        torch.save({"news_vecs": self.news_vecs.cpu(), "version": self.version}, tmp)
        os.replace(tmp, path)

    def load_table(self, path: str) -> "Recommender":
        blob = torch.load(path, map_location="cpu", weights_only=True)
        self.refresh(blob["news_vecs"])
        self.version = int(blob.get("version", self.version))
        return self

    def _need_table(self) -> None:
        if self.news_vecs is None:
            self.refresh()

    def pad_histories(self, histories: Sequence[Sequence[int]]) -> torch.Tensor:
        """Lists of news rows → ``[B, L]`` int64, the most recent ``L`` clicks, 0-padded on the right like training."""
        out = torch.zeros(len(histories), self.L, dtype=torch.long)
        for i, h in enumerate(histories):
            h = list(h)[-self.L:]
            if h:
                out[i, : len(h)] = torch.as_tensor(h, dtype=torch.long)
        return out

    # ---- requests ---------------------------------------------------------------------------------------------
    @torch.no_grad()
    def user_vectors(self, histories) -> torch.Tensor:
        """``[B, L]`` news rows (0 = padding) → ``[B, D]`` user vectors (dropout off)."""
        self._need_table()
        his = _as_ids(histories, self.device)
        ue = self.model.user_encoder
        was = ue.training
        ue.eval()
        try:
            mask = (his != 0) if self.model.cfg.mask_padding else None
            return ue(self.news_vecs[his], mask)
        finally:
            ue.train(was)

    @torch.no_grad()
    def score(self, histories, candidates) -> torch.Tensor:
        """``[B, L]`` histories, ``[B, C]`` candidate rows → ``[B, C]`` click scores (what ``forward`` returns)."""
        u = self.user_vectors(histories)
        cand = _as_ids(candidates, self.device)
        return torch.sigmoid(torch.bmm(self.news_vecs[cand], u.unsqueeze(-1)).squeeze(-1))

    @torch.no_grad()
    def recommend(self, histories, k: int = 10, exclude_seen: bool = True,
                  allowed: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Top-``k`` news rows and their scores for every history: ``([B, k] int64, [B, k] fp32)``.

        Row 0 (``<unk>`` / padding) is never returned; ``exclude_seen`` drops what the user already clicked; ``allowed``
        is an optional ``[N]`` bool mask of servable rows (fresh, not withdrawn, ...)."""
        self._need_table()
        his = _as_ids(histories, self.device)
        u = self.user_vectors(his)
        logits = (u.to(self.table_dtype) @ self._table.t()).float()            # [B, N]
        neg = torch.finfo(torch.float32).min
        logits[:, 0] = neg
        if allowed is not None:
            logits.masked_fill_(~allowed.to(self.device).bool().view(1, -1), neg)
        if exclude_seen:
            logits.scatter_(1, his, neg)                                       # histories hold row numbers; 0 is already out
        k = min(k, logits.shape[1] - 1)
        top, idx = torch.topk(logits, k, dim=1)
        return idx, torch.sigmoid(top)
