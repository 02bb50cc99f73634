"""The sm_100a step pipeline: everything between "ids are on the device" and "gradients are in
the flat buffer" runs on this framework's own kernels.

Pipeline (reference call sites in parentheses):

  dedup ids                                  (none upstream: model.py:41-61 encodes every duplicate)
  embed + LayerNorm                          (HF Embeddings via encoder.py:27)
  6 x [ QKV GEMM · attention · out-proj GEMM(+residual) · LN · FFN1 GEMM(+GELU) · FFN2 GEMM(+residual) · LN ]
  text head: fc1 GEMM(+tanh) · additive pooling · fc GEMM          (encoder.py:28-29)
  user encoder + scoring + soft-max CE, forward and backward      (encoder.py:49-56, model.py:121-126)
  news-gradient segment-sum                  (main.py:20-42)
  text-head backward with the injected per-news gradient          (model.py:72-90)

All GEMMs are ``fr_gemm_bf16`` (tcgen05 + TMEM + TMA). Buffers are allocated once for the
worst case (every id distinct) and the de-dup count stays on the device, so the whole step is
free of host synchronisation and can be captured in a CUDA graph.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn.functional as F

from .. import ops
from ..ops import dispatch as D
from ..config import FedRecConfig


class PackedTrunk:
    """bf16 copies of the frozen DistilBERT weights in the layout the kernels want."""

    def __init__(self, bert: torch.nn.Module):
        self.bert = bert
        self.repack()

    @torch.no_grad()
    def repack(self) -> None:
        b = self.bert
        bf = torch.bfloat16
        self.word = D.cast_bf16(b.embeddings.word_embeddings.weight.data)
        self.pos = D.cast_bf16(b.embeddings.position_embeddings.weight.data)
        self.emb_g = b.embeddings.LayerNorm.weight.data.float().contiguous()
        self.emb_b = b.embeddings.LayerNorm.bias.data.float().contiguous()
        self.layers = []
        for blk in b.transformer.layer:
            a = blk.attention
            wqkv = torch.cat([a.q_lin.weight.data, a.k_lin.weight.data, a.v_lin.weight.data], 0)
            self.layers.append({
                "wqkv": D.cast_bf16(wqkv),
                "bqkv": torch.cat([a.q_lin.bias.data, a.k_lin.bias.data, a.v_lin.bias.data]).float().contiguous(),
                "wo": D.cast_bf16(a.out_lin.weight.data), "bo": a.out_lin.bias.data.float().contiguous(),
                "ln1_g": blk.sa_layer_norm.weight.data.float().contiguous(),
                "ln1_b": blk.sa_layer_norm.bias.data.float().contiguous(),
                "w1": D.cast_bf16(blk.ffn.lin1.weight.data), "b1": blk.ffn.lin1.bias.data.float().contiguous(),
                "w2": D.cast_bf16(blk.ffn.lin2.weight.data), "b2": blk.ffn.lin2.bias.data.float().contiguous(),
                "ln2_g": blk.output_layer_norm.weight.data.float().contiguous(),
                "ln2_b": blk.output_layer_norm.bias.data.float().contiguous(),
            })
        self.heads = b.config.n_heads
        self.dim = b.config.dim
        self.ffn = b.config.hidden_dim


class FusedStep:
    def __init__(self, model, cfg: FedRecConfig):
        ops.require_native()
        self.model, self.cfg = model, cfg
        self.dev = model.device
        assert self.dev.type == "cuda"
        self.trunk = PackedTrunk(model.text_encoder.DistillBert)
        self._bufs: Dict[int, Dict[str, torch.Tensor]] = {}
        self._dedup_ws: Optional[torch.Tensor] = None
        self._gen = torch.Generator(device=self.dev)
        self._gen.manual_seed(cfg.seed + 17)

    def repack_trunk(self) -> None:
        self.trunk.repack()

    # ---- buffers ---------------------------------------------------------------------------------
    def _buffers(self, K: int) -> Dict[str, torch.Tensor]:
        if K in self._bufs:
            return self._bufs[K]
        T = self.model.news_tokens.shape[1]
        R = K * T
        dim, ffn = self.trunk.dim, self.trunk.ffn
        bf, dev = torch.bfloat16, self.dev
        z = lambda *s, dt=bf: torch.zeros(*s, device=dev, dtype=dt)  # noqa: E731
        hq = self.model.text_encoder.additive_attention.att_fc1.out_features
        b = {"x": z(R, dim), "y": z(R, dim), "ctx": z(R, dim), "qkv": z(R, 3 * dim), "ffn": z(R, ffn),
             "e": z(R, hq), "pooled": z(K, dim), "alpha": z(R, dt=torch.float32),
             "news_vecs": z(K, self.cfg.news_dim, dt=torch.float32)}
        self._bufs[K] = b
        return b

    # ---- trunk + head forward ------------------------------------------------------------------------
    @torch.no_grad()
    def encode(self, uniq: torch.Tensor, count: Optional[torch.Tensor], K: int) -> Dict[str, torch.Tensor]:
        """``uniq`` int32 [K] news rows (first ``*count`` valid) → buffers with ``hidden`` (=x) and ``news_vecs``."""
        m, tr, b = self.model, self.trunk, self._buffers(K)
        T = m.news_tokens.shape[1]
        x, y, ctx, qkv, ffn = b["x"], b["y"], b["ctx"], b["qkv"], b["ffn"]
        D.bert_embed_ln(m.news_tokens, uniq, count, K, T, tr.word, tr.pos, tr.emb_g, tr.emb_b, x)
        for L in tr.layers:
            D.linear_bf16(x, L["wqkv"], L["bqkv"], "bias", out=qkv, m_dev=count, m_scale=T)
            D.bert_attention(qkv, m.news_mask, uniq, count, K, T, tr.heads, ctx)
            D.linear_bf16(ctx, L["wo"], L["bo"], "resid", resid=x, out=y, m_dev=count, m_scale=T)
            D.layernorm_bf16(y, L["ln1_g"], L["ln1_b"], out=x, count=count, rows_per_item=T)
            D.linear_bf16(x, L["w1"], L["b1"], "gelu", out=ffn, m_dev=count, m_scale=T)
            D.linear_bf16(ffn, L["w2"], L["b2"], "resid", resid=x, out=y, m_dev=count, m_scale=T)
            D.layernorm_bf16(y, L["ln2_g"], L["ln2_b"], out=x, count=count, rows_per_item=T)
        te = m.text_encoder
        aa = te.additive_attention
        w1 = D.cast_bf16(aa.att_fc1.weight.data)
        D.linear_bf16(x, w1, aa.att_fc1.bias.data, "tanh", out=b["e"], m_dev=count, m_scale=T)
        mask = m.news_mask if te.mask_padding else None
        D.text_pool(x, b["e"], aa.att_fc2.weight.data.view(-1), aa.att_fc2.bias.data, mask, uniq, count, K, T,
                    b["pooled"], b["alpha"])
        wfc = D.cast_bf16(te.fc.weight.data)
        D.linear_bf16(b["pooled"], wfc, te.fc.bias.data, "bias", out=b["news_vecs"], m_dev=count, m_scale=1)
        return b

    def _ids(self, cand, his):
        ids = torch.cat([his.reshape(-1), cand.reshape(-1)]).to(torch.int32)
        K = ids.numel()
        if self.cfg.dedup:
            need = ops.require_native().fr_dedup_workspace_ints(self.model.news_tokens.shape[0])
            if self._dedup_ws is None or self._dedup_ws.numel() < need:
                self._dedup_ws = torch.empty(need, dtype=torch.int32, device=self.dev)
            uniq, count, inv = D.dedup(ids, self.model.news_tokens.shape[0], self._dedup_ws)
        else:
            uniq, count, inv = ids, None, torch.arange(K, device=self.dev, dtype=torch.int32)
        return ids, uniq, count, inv, K

    # ---- full step -----------------------------------------------------------------------------------------
    def run(self, cand: torch.Tensor, his: torch.Tensor, train: bool = True) -> torch.Tensor:
        m, cfg = self.model, self.cfg
        B, C = cand.shape
        L = his.shape[1]
        ids, uniq, count, inv, K = self._ids(cand, his)
        b = self.encode(uniq, count, K)
        m.train(train)
        # --- user encoder + loss (autograd over fp32 news vectors) ---
        nv = b["news_vecs"].detach().requires_grad_(train)
        inv_l = inv.long()
        his_vec = nv[inv_l[: B * L]].view(B, L, -1)
        cand_vec = nv[inv_l[B * L:]].view(B, C, -1)
        if train and cfg.news_grad_noise > 0:
            std = cfg.news_grad_noise

            def _noise(g):
                return g + torch.randn(g.shape, device=g.device, dtype=g.dtype, generator=self._gen) * std
            his_vec.register_hook(_noise)
            cand_vec.register_hook(_noise)
        his_mask = (his != 0) if cfg.mask_padding else None
        user = m.user_encoder(his_vec, his_mask)
        loss, _ = ops.score_ce(cand_vec, user)
        if not train:
            return loss.detach()
        m.flat.zero_grad()
        loss.backward()
        # --- text head backward with the per-news gradient injected (head-only; trunk is frozen) ---
        T = m.news_tokens.shape[1]
        hidden = b["x"].view(K, T, -1).float()
        te = m.text_encoder
        msk = m.news_mask[uniq.long()] if te.mask_padding else None
        with torch.enable_grad():
            vec = te.head(hidden, msk)
            vec.backward(nv.grad)
        return loss.detach()

    @torch.no_grad()
    def scores(self, cand: torch.Tensor, his: torch.Tensor) -> torch.Tensor:
        m = self.model
        B, C = cand.shape
        L = his.shape[1]
        ids, uniq, count, inv, K = self._ids(cand, his)
        b = self.encode(uniq, count, K)
        nv = b["news_vecs"]
        inv_l = inv.long()
        was = m.training
        m.eval()
        his_mask = (his != 0) if self.cfg.mask_padding else None
        user = m.user_encoder(nv[inv_l[: B * L]].view(B, L, -1), his_mask)
        m.train(was)
        return torch.sigmoid(torch.bmm(nv[inv_l[B * L:]].view(B, C, -1), user.unsqueeze(-1)).squeeze(-1))
