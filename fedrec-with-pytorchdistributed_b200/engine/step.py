"""One local training step: batch of impressions → gradients in the flat buffer.

Reference flow being replaced: ``UserModel.forward`` (2·B DistilBERT
calls through numpy), ``loss.backward()``, leaf grads to CPU, ``process_news_grad``
Python loop, ``collect``, then ``update_news_grad`` re-running DistilBERT with grad.

Two implementations with identical results up to dtype:

``torch``  end-to-end autograd in fp32: unique ids → frozen trunk (no_grad) → head →
           gather → user encoder → score → CE → backward. This is the in-repo oracle
           and the CPU path. Differentiating through the gather *is* the segment-sum
           of news gradients followed by the injected head backward.
``fused``  the sm_100a kernel pipeline in ``engine/fused.py`` (bf16 tensor-core trunk
           and head, fused user-encoder forward/backward, device-side de-dup).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from ..config import FedRecConfig


class TrainStep:
    def __init__(self, model, cfg: Optional[FedRecConfig] = None, impl: Optional[str] = None):
        self.model = model
        self.cfg = cfg or model.cfg
        dev = model.device
        if impl is None:
            impl = "fused" if (dev.type == "cuda" and self.cfg.fused) else "torch"
            if impl == "fused" and self.cfg.dtype == "fp32":
                impl = "torch"               # dtype=fp32: the fp32 PyTorch step, for debugging numerics against the bf16 kernels
            if impl == "fused" and self.cfg.trunk_dropout_in_update:
                import warnings              # model.py:73 (DistilBERT in train() for the gradient pass) exists only in the oracle step
                warnings.warn("fedrec_b200: trunk_dropout_in_update is honoured by the PyTorch step only (the sm_100a "
                              "pipeline reuses the eval-mode trunk output); running the PyTorch path", RuntimeWarning)
                impl = "torch"
            if impl == "fused":
                from .fused import fused_unsupported_reason
                why = fused_unsupported_reason(model, self.cfg)
                if why is not None:          # outside the shapes the kernels are built for: say so, use the oracle path
                    import warnings
                    warnings.warn(f"fedrec_b200: sm_100a step pipeline not available for this model ({why}); "
                                  f"running the PyTorch path", RuntimeWarning)
                    impl = "torch"
        self.impl = impl
        self._fused = None
        if impl == "fused":
            from .fused import FusedStep
            self._fused = FusedStep(model, self.cfg)
            model._fused = self._fused
        self._gen = torch.Generator(device=dev)
        self._gen.manual_seed(self.cfg.seed + 17)

    # ---------------------------------------------------------------------------------
    def run(self, cand: torch.Tensor, his: torch.Tensor, train: bool = True) -> torch.Tensor:
        """``cand [B,C]``, ``his [B,L]`` integer news rows on the model device → scalar loss tensor.

        Leaves d(loss)/d(trainable) in ``model.flat.grad`` (overwritten, not accumulated).
        """
        if self._fused is not None:
            return self._fused.run(cand, his, train=train)
        return self._run_torch(cand, his, train)

    # ---- optimiser pipelined into the step (fused implementation only) -------------------------------------------
    def attach_optimizer(self, grad_agg, lr: float, b1: float, b2: float, eps: float) -> bool:
        """Let ``run`` also apply the update (aggregation + Adam of step i at the start of step i+1, beside the frozen
        trunk; ``flush`` applies the last one). Returns False when this step implementation cannot pipeline — the
        caller then runs aggregation and Adam itself after every ``run``."""
        if self._fused is None or not grad_agg.on_device_chain:
            return False
        self._fused.attach_optimizer(grad_agg, lr, b1, b2, eps)
        return True

    @property
    def pipelined(self) -> bool:
        return self._fused is not None and self._fused._opt is not None

    def flush(self) -> None:
        if self._fused is not None:
            self._fused.flush()

    def detach_optimizer(self) -> None:
        if self._fused is not None:
            self._fused.detach_optimizer()

    def run_packed(self, ids: torch.Tensor, B: int, C: int, L: int) -> torch.Tensor:
        """``run`` with the ids already in one int32 buffer ``[B*L + B*C]`` (histories first), device or pinned host."""
        if self._fused is not None:
            return self._fused.run_packed(ids, B, C, L)
        d = ids.to(self.model.device)
        return self._run_torch(d[B * L:].view(B, C), d[: B * L].view(B, L), True)

    # ---- end-to-end stepping from host memory ------------------------------------------------------------------
    def enable_host_io(self, B: int, C: int, L: int, slots: int = 4) -> None:
        """Prepare ``run_host`` / ``loss_of`` for batches of this shape (fused: pinned host-mapped rings read and written
        by the step's own first and last kernel; PyTorch path: plain copies)."""
        self._io_shape = (B, C, L)
        self._io_results = {}
        self._io_issued = 0
        if self._fused is not None:
            self._fused.enable_host_io(B * (L + C), slots)

    def run_host(self, ids) -> int:
        """One training step on ids held in host memory (int32 ``[B*L + B*C]``, histories first). Returns a ticket;
        ``loss_of(ticket)`` delivers that step's loss to the host. The host may run a few steps ahead of the results."""
        B, C, L = self._io_shape
        if self._fused is not None:
            return self._fused.run_host(ids, B, C, L)
        d = torch.as_tensor(ids).to(self.model.device)
        self._io_issued += 1
        self._io_results[self._io_issued] = self._run_torch(d[B * L:].view(B, C), d[: B * L].view(B, L), True)
        return self._io_issued

    def loss_of(self, ticket: int) -> float:
        if self._fused is not None:
            return self._fused.loss_of(ticket)
        return float(self._io_results.pop(ticket))

    def scores(self, cand: torch.Tensor, his: torch.Tensor) -> torch.Tensor:
        """Inference: ``sigmoid(cand · user)`` ``[B, C]`` without touching gradients."""
        if self._fused is not None:
            return self._fused.scores(cand, his)
        m = self.model
        was = m.training
        m.eval()
        with torch.no_grad():
            ids = torch.cat([his.reshape(-1), cand.reshape(-1)]).long()
            uniq, inv = torch.unique(ids, return_inverse=True)
            if self.cfg.precomputed_news and m.news_features is not None:
                vec = m.text_encoder.fc(m.news_features[uniq])
            else:
                vec = m.encode_unique(uniq, keep_cache=False, use_fused=False)
            B, L = his.shape
            his_mask = (his != 0) if self.cfg.mask_padding else None
            user = m.user_encoder(vec[inv[: B * L]].view(B, L, -1), his_mask)
            score = torch.sigmoid(torch.bmm(vec[inv[B * L:]].view(B, cand.shape[1], -1), user.unsqueeze(-1)).squeeze(-1))
        m.train(was)
        return score

    # ---------------------------------------------------------------------------------
    def _run_torch(self, cand, his, train):
        m, cfg = self.model, self.cfg
        B, C = cand.shape
        L = his.shape[1]
        m.train(train)
        ids = torch.cat([his.reshape(-1), cand.reshape(-1)]).long()
        if cfg.dedup:
            uniq, inv = torch.unique(ids, return_inverse=True)
        else:
            uniq, inv = ids, torch.arange(ids.numel(), device=ids.device)
        if cfg.precomputed_news:
            # BASELINE.json config 1: no trunk in the step; the 768-d features are a cached table, the 768->400 projection
            # and the user encoder train (the text head's additive attention is frozen into the cache)
            if m.news_features is None:
                m.build_news_features()
            vec = m.text_encoder.fc(m.news_features[uniq])
        elif cfg.trunk_dropout_in_update and train:
            # upstream quirk (model.py:73): the gradient pass re-runs DistilBERT in train mode
            m.text_encoder.DistillBert.train()
            with torch.no_grad():
                hidden = m.text_encoder.trunk(m.news_tokens[uniq].long(), m.news_mask[uniq])
            vec = m.text_encoder.head(hidden, m.news_mask[uniq])
        else:
            vec = m.encode_unique(uniq, keep_cache=False, with_grad=True, use_fused=False)   # the oracle stays pure PyTorch
        his_vec = vec[inv[: B * L]].view(B, L, -1)
        cand_vec = vec[inv[B * L:]].view(B, C, -1)
        if train and cfg.news_grad_noise > 0:
            std = cfg.news_grad_noise

            def _noise(g):
                return g + torch.randn(g.shape, device=g.device, dtype=g.dtype, generator=self._gen) * std
            his_vec.register_hook(_noise)
            cand_vec.register_hook(_noise)
        his_mask = (his != 0) if cfg.mask_padding else None
        user = m.user_encoder(his_vec, his_mask)
        score = torch.sigmoid(torch.bmm(cand_vec, user.unsqueeze(-1)).squeeze(-1))
        loss = F.cross_entropy(score, torch.zeros(B, dtype=torch.long, device=score.device))
        if train:
            m.flat.zero_grad()
            loss.backward()
        return loss.detach()
