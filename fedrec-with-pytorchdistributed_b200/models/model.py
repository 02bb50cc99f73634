"""Top-level recommendation model.

Behavioural spec: reference model.py — UserModel :10-129 (gen_news_vecs :41-61,
update :66-70, update_news_grad :72-90, update_user_grad :92-94, collect :97-109,
forward :111-129) and ServerUserModel :134-248.  The public names and call
signatures are kept; what sits behind them is rebuilt for one-GPU-per-client:

* the news token table lives on the device, so "encode these ids" never touches
  the host (the reference round-trips through numpy four times per batch,
  SURVEY.md §3.1);
* ids are de-duplicated per step and the frozen trunk runs once per distinct news
  item (exact, because DistilBERT is frozen and in eval mode);
* trunk activations of the step are cached, so the text-head update
  (``update_news_grad``) is a head-only forward/backward with the accumulated
  per-news gradient injected — no second DistilBERT pass;
* the 16 trainable tensors are views into one flat buffer (``utils/flat.py``) that
  the aggregation kernels and the fused Adam operate on.

Reference defects that are *not* reproduced: gradient doubling in
``collect`` (§9.2), CPU labels (§9.5), ``ServerUserModel`` failing to construct
(§9.4).
"""
from __future__ import annotations

import os
import time

from typing import Dict, List, Optional, Tuple, Union

import numpy as np
import torch
from torch import nn

from ..config import FedRecConfig
from ..utils.flat import FlatState
from .distilbert import DistilBertConfigLite
from .encoder import TextEncoder, UserEncoder

IdLike = Union[torch.Tensor, np.ndarray, List]


def _as_ids(x: IdLike, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.long, non_blocking=True)
    return torch.as_tensor(np.asarray(x), dtype=torch.long).to(device)


class UserModel(nn.Module):
    def __init__(self, news_dataset=None, news_index=None, device=None, news_embedding_dim: int = 400,
                 num_attention_heads: int = 20, query_vector_dim: int = 200,
                 config: Optional[FedRecConfig] = None, bert_config: Optional[DistilBertConfigLite] = None,
                 flat_allocator=None):
        super().__init__()
        self.cfg = config or FedRecConfig()
        self.device = torch.device(device) if device is not None else self.cfg.resolved_device()
        # cfg.bert_path: a local HuggingFace-format DistilBERT directory (upstream downloads the pretrained weights,
        # encoder.py:19; here they are read from disk when given, else the trunk is random-init from the config)
        pretrained = self.cfg.bert_path if (bert_config is None and self.cfg.bert_path) else None
        if pretrained is not None and not os.path.isdir(pretrained):
            raise FileNotFoundError(f"bert_path={pretrained!r} is not a directory")
        if bert_config is None and pretrained is None:
            bert_config = DistilBertConfigLite(
                vocab_size=self.cfg.vocab_size, max_position_embeddings=self.cfg.max_positions,
                n_layers=self.cfg.bert_layers, n_heads=self.cfg.bert_heads, dim=self.cfg.bert_dim,
                hidden_dim=self.cfg.bert_ffn, hidden_size=self.cfg.bert_dim)
        self.text_encoder = TextEncoder(news_embedding_dim, bert_config=bert_config, pretrained=pretrained or "",
                                        mask_padding=self.cfg.mask_padding).to(self.device)
        self.user_encoder = UserEncoder(news_embedding_dim, num_attention_heads, query_vector_dim,
                                        head_dim=news_embedding_dim // num_attention_heads,
                                        dropout_rate=self.cfg.dropout,
                                        mask_padding=self.cfg.mask_padding).to(self.device)
        for p in self.text_encoder.DistillBert.parameters():      # model.py:25-26
            p.requires_grad = False
        self.news_dataset = news_dataset                          # kept for signature parity (unused upstream too)
        self.criterion = nn.CrossEntropyLoss()
        self.cnt = 0
        self.time = time.time()                                   # model.py:31 (never read upstream either)
        self.set_news_index(news_index)
        trainable = [(n, p) for n, p in self.named_parameters() if p.requires_grad]
        self.flat = FlatState(trainable, self.device, allocator=flat_allocator)
        # With average_frozen the 100 frozen DistilBERT tensors also become views of one flat buffer (in the peer-mapped
        # heap when there is one), so the full 270 MB state_dict is averaged/broadcast by one kernel (upstream: 116
        # all-reduces per epoch, Parameter_Averaging_main.py:144-148).
        self.flat_frozen = None
        if self.cfg.average_frozen:
            frozen = [(n, p) for n, p in self.named_parameters() if not p.requires_grad]
            self.flat_frozen = FlatState(frozen, self.device, allocator=flat_allocator, with_adam=False, with_grad=False)
        self._init_grad_param()
        self._cache: Optional[Dict[str, torch.Tensor]] = None
        self._fused = None        # engine.fused.FusedStep, attached lazily by the engine
        self.news_features: Optional[torch.Tensor] = None     # [N, 768] cache for cfg.precomputed_news
        self._news_features_bf16: Optional[torch.Tensor] = None

    # ---- news table -------------------------------------------------------------
    def set_news_index(self, news_index) -> None:
        self.news_index = news_index
        if news_index is None:
            self.news_tokens = None
            self.news_mask = None
            return
        ni = torch.as_tensor(np.asarray(news_index))
        self.news_tokens = ni[:, 0, :].to(self.device, torch.int32).contiguous()
        self.news_mask = ni[:, 1, :].to(self.device, torch.int32).contiguous()

    def on_weights_changed(self) -> None:
        """Call after the frozen trunk's weights are replaced (load/broadcast) to refresh packed copies."""
        if self._fused is not None:
            self._fused.repack_trunk()

    # ---- optimiser ---------------------------------------------------------------
    def _init_grad_param(self) -> None:
        self._ng_ids: List[torch.Tensor] = []
        self._ng_vals: List[torch.Tensor] = []
        self.flat.zero_grad()

    @property
    def news_grads(self) -> Dict[int, torch.Tensor]:
        """Accumulated per-news gradients as ``{row: grad[400]}`` (materialised on demand)."""
        if not self._ng_ids:
            return {}
        ids, vals = self._reduced_news_grads()
        return {int(i): v for i, v in zip(ids.tolist(), vals)}

    def sync_optimizer(self) -> None:
        """Apply an update the pipelined step still owes (engine/fused.py ``attach_optimizer``); no-op otherwise."""
        if self._fused is not None:
            self._fused.flush()

    def state_dict(self, *args, **kwargs):
        self.sync_optimizer()
        return super().state_dict(*args, **kwargs)

    def adam_step(self) -> None:
        from .. import ops
        self.sync_optimizer()
        ops.adam_flat(self.flat, self.cfg.lr, self.cfg.beta1, self.cfg.beta2, self.cfg.adam_eps)

    # ---- encoding ------------------------------------------------------------------
    def _fused_engine(self):
        """The sm_100a pipeline, created on first use for CUDA models (engine/fused.py)."""
        if self._fused is None and self.device.type == "cuda" and self.cfg.fused:
            from ..engine.fused import FusedStep, fused_unsupported_reason
            if fused_unsupported_reason(self, self.cfg) is None:
                self._fused = FusedStep(self, self.cfg)
        return self._fused

    def encode_unique(self, uniq: torch.Tensor, keep_cache: bool = True, with_grad: bool = False,
                      use_fused: Optional[bool] = None) -> torch.Tensor:
        """``uniq`` int64 ``[U]`` news rows → ``[U, D]`` news vectors (trunk under no_grad/eval).

        On CUDA the frozen trunk runs on the framework's own kernels (bf16 tcgen05 GEMMs, tensor-core attention); the
        trainable head runs there too unless ``with_grad`` asks for an autograd graph through it."""
        te = self.text_encoder
        fused = self._fused_engine() if use_fused in (None, True) else None
        if fused is not None:
            U = uniq.numel()
            cap = (U + 255) // 256 * 256                 # bucketed capacity: a bounded set of activation buffers
            padded = torch.zeros(cap, dtype=torch.int32, device=self.device)
            padded[:U] = uniq
            b = fused.encode(padded, torch.tensor([U], dtype=torch.int32, device=self.device), cap)
            T = self.news_tokens.shape[1]
            hidden = b["x"][: U * T].view(U, T, -1).float()
            msk = self.news_mask[uniq.long()]
            if keep_cache:
                self._cache = {"uniq": uniq, "hidden": hidden, "mask": msk}
            if with_grad:
                return te.head(hidden, msk)
            return b["news_vecs"][:U].clone()
        was_training = te.DistillBert.training
        te.DistillBert.eval()
        with torch.no_grad():
            tok = self.news_tokens[uniq].long()
            msk = self.news_mask[uniq]
            hidden = te.trunk(tok, msk)
        te.DistillBert.train(was_training)
        if keep_cache:
            self._cache = {"uniq": uniq, "hidden": hidden, "mask": msk}
        if with_grad:
            return te.head(hidden, msk)
        with torch.no_grad():
            return te.head(hidden, msk)

    def get_news_vecs(self, idx) -> torch.Tensor:
        """Vectors of news rows ``idx`` (upstream ``get_news_vecs`` reads an attribute it never sets, model.py:63-64)."""
        return self.gen_news_vecs(idx)

    def gen_news_vecs(self, nids: IdLike) -> torch.Tensor:
        """News vectors for an id array of any shape ``[...]`` → ``[..., D]`` (no grad).

        The reference loops over rows and returns a list of numpy arrays
        (model.py:41-61); here it is one de-duplicated device pass.
        """
        ids = _as_ids(nids, self.device)
        uniq, inv = torch.unique(ids.reshape(-1), return_inverse=True)
        vec = self.encode_unique(uniq)
        if torch.isnan(vec).any():
            raise ValueError("news_vecs contains nan")
        return vec[inv].view(*ids.shape, -1)

    # ---- precomputed news features (BASELINE.json config 1) -------------------------------------------------------
    @torch.no_grad()
    def build_news_features(self, batch: int = 2048) -> torch.Tensor:
        """Cache ``[N, 768]`` pooled DistilBERT features of the whole news table (frozen trunk + the text head's additive
        attention *as it is now*), so that ``cfg.precomputed_news`` steps train the user encoder and the 768→400
        projection without ever running the trunk. This is the training-side counterpart of what upstream's
        ``UserDataset`` (dataset.py:92-109) and ``evaluation_split`` (evaluation_functions.py:33-46) anticipate:
        news vectors computed once, users encoded over the cache."""
        self.sync_optimizer()
        te = self.text_encoder
        N = self.news_tokens.shape[0]
        feats = torch.empty(N, te.fc.in_features, device=self.device, dtype=torch.float32)
        fused = self._fused_engine()
        for i in range(0, N, batch):
            ids = torch.arange(i, min(N, i + batch), device=self.device)
            if fused is not None:
                U = ids.numel()
                cap = (U + 255) // 256 * 256
                padded = torch.zeros(cap, dtype=torch.int32, device=self.device)
                padded[:U] = ids
                b = fused.encode(padded, torch.tensor([U], dtype=torch.int32, device=self.device), cap)
                feats[i:i + U] = b["pooled"][:U].float()
            else:
                was = te.DistillBert.training
                te.DistillBert.eval()
                hidden = te.trunk(self.news_tokens[ids].long(), self.news_mask[ids])
                te.DistillBert.train(was)
                feats[i:i + ids.numel()] = te.additive_attention(hidden, self.news_mask[ids] if te.mask_padding else None)
        self.set_news_features(feats)
        return feats

    def set_news_features(self, feats: torch.Tensor) -> None:
        """Install a ``[N, 768]`` feature table (computed elsewhere, e.g. by a server with the full trunk)."""
        assert feats.dim() == 2 and feats.shape[1] == self.text_encoder.fc.in_features
        self.news_features = feats.to(self.device, torch.float32).contiguous()
        self._news_features_bf16 = None

    def news_features_bf16(self) -> torch.Tensor:
        if getattr(self, "_news_features_bf16", None) is None:
            self._news_features_bf16 = self.news_features.to(torch.bfloat16).contiguous()
        return self._news_features_bf16

    # ---- precomputed-vector inference (the path upstream's UserDataset / evaluation_split anticipate) ---------------
    @torch.no_grad()
    def encode_all_news(self, batch: int = 2048) -> torch.Tensor:
        """News vectors of the whole table ``[N, D]`` (fp32, on the model device), ``batch`` rows per pass."""
        N = self.news_tokens.shape[0]
        out = torch.empty(N, self.cfg.news_dim, device=self.device)
        for i in range(0, N, batch):
            ids = torch.arange(i, min(N, i + batch), device=self.device)
            out[i:i + ids.numel()] = self.encode_unique(ids, keep_cache=False)
        return out

    @torch.no_grad()
    def encode_users(self, samples, nid2index, news_vecs: torch.Tensor, batch: int = 1024) -> torch.Tensor:
        """User vectors ``[S, D]`` from precomputed ``news_vecs`` for sample tuples ``(_, _, _, history, _)``."""
        L = self.cfg.max_his_len
        his = torch.zeros(len(samples), L, dtype=torch.long)
        for s, sam in enumerate(samples):
            h = [nid2index[n] for n in sam[3]][-L:]
            his[s, : len(h)] = torch.tensor(h, dtype=torch.long)
        his = his.to(self.device)
        was = self.user_encoder.training
        self.user_encoder.eval()
        outs = []
        for i in range(0, len(samples), batch):
            hb = his[i:i + batch]
            mask = (hb != 0) if self.cfg.mask_padding else None
            outs.append(self.user_encoder(news_vecs[hb], mask))
        self.user_encoder.train(was)
        return torch.cat(outs) if outs else torch.empty(0, self.cfg.news_dim, device=self.device)

    # ---- forward / backward plumbing (reference-style API) -----------------------------
    def forward(self, candidate_news: IdLike, his_news: IdLike, labels: Optional[IdLike] = None,
                compute_loss: bool = True):
        cand = _as_ids(candidate_news, self.device)
        his = _as_ids(his_news, self.device)
        B = cand.shape[0]
        ids = torch.cat([his.reshape(-1), cand.reshape(-1)])
        uniq, inv = torch.unique(ids, return_inverse=True)
        vec = self.encode_unique(uniq)
        nh = his.numel()
        his_vec = vec[inv[:nh]].view(B, his.shape[1], -1).detach().requires_grad_(True)
        candidate_vecs = vec[inv[nh:]].view(B, cand.shape[1], -1).detach().requires_grad_(True)
        his_mask = (his != 0) if self.cfg.mask_padding else None
        user_vector = self.user_encoder(his_vec, his_mask)
        score = torch.sigmoid(torch.bmm(candidate_vecs, user_vector.unsqueeze(-1)).squeeze(-1))
        if not compute_loss:
            return score, candidate_vecs, his_vec
        if labels is None:
            labels = torch.zeros(B, dtype=torch.long, device=self.device)
        labels = _as_ids(labels, self.device).view(-1)
        loss = self.criterion(score, labels)
        return loss, score, candidate_vecs, his_vec

    def collect(self, news_grad, user_grad=None) -> None:
        """Accumulate per-news gradients (and, if given, user-encoder gradients).

        ``news_grad`` is either ``{row: grad}`` (reference layout) or an
        ``(ids[K], grads[K, D])`` pair.  ``user_grad`` entries that alias the live
        ``.grad`` tensors are skipped, which removes the reference's doubling.
        """
        if isinstance(news_grad, dict):
            if news_grad:
                keys = list(news_grad.keys())
                ids = torch.as_tensor(np.asarray(keys, dtype=np.int64), device=self.device)
                vals = torch.stack([torch.as_tensor(news_grad[k]) for k in keys]).to(self.device, torch.float32)
                self._ng_ids.append(ids)
                self._ng_vals.append(vals)
        elif news_grad is not None:
            ids, vals = news_grad
            self._ng_ids.append(_as_ids(ids, self.device).reshape(-1))
            self._ng_vals.append(vals.to(self.device, torch.float32).reshape(-1, vals.shape[-1]))
        if user_grad:
            for name, p in self.user_encoder.named_parameters():
                g = user_grad.get(name)
                if g is None or p.grad is None:
                    continue
                if g.data_ptr() == p.grad.data_ptr():
                    continue
                p.grad += g.to(p.grad.device)

    def _reduced_news_grads(self) -> Tuple[torch.Tensor, torch.Tensor]:
        from .. import ops
        ids = torch.cat(self._ng_ids)
        vals = torch.cat(self._ng_vals)
        uniq, inv = torch.unique(ids, return_inverse=True)
        return uniq, ops.segment_sum(vals, inv, uniq.numel())

    def update_user_grad(self) -> None:
        """Kept for API parity; the flat Adam in ``update`` covers both parameter groups."""

    def update_news_grad(self) -> None:
        """Head-only forward/backward with the accumulated news gradients injected (model.py:72-90)."""
        if not self._ng_ids:
            return
        uniq, grads = self._reduced_news_grads()
        cache = self._cache
        if cache is not None and cache["uniq"].shape == uniq.shape and bool((cache["uniq"] == uniq).all()):
            hidden, msk = cache["hidden"], cache["mask"]
        else:
            self.encode_unique(uniq)
            hidden, msk = self._cache["hidden"], self._cache["mask"]
        with torch.enable_grad():
            vec = self.text_encoder.head(hidden, msk)
            vec.backward(grads.to(vec.dtype))

    def update(self) -> None:
        self.update_news_grad()          # fills text-head grads in the flat buffer
        self.adam_step()                 # one Adam step over all 16 trainable tensors
        self._init_grad_param()
        self.cnt += 1

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        self.sync_optimizer()
        out = super().load_state_dict(state_dict, strict=strict, assign=False)
        self.flat.rebind()
        if self.flat_frozen is not None:
            self.flat_frozen.rebind()
        self.on_weights_changed()
        return out


class ServerUserModel(UserModel):
    """The aggregation-only role: same parameters, no data (server.py:66)."""

    def __init__(self, device=None, config: Optional[FedRecConfig] = None, **kw):
        super().__init__(news_dataset=None, news_index=None, device=device, config=config, **kw)
