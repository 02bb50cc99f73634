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

All GEMMs are ``fr_gemm_bf16`` / ``fr_gemm_bf16_fold`` (tcgen05 + TMEM + TMA); with ``cfg.ln_fold`` 11 of the trunk's 12
LayerNorms live in the epilogues of the GEMMs around them. Buffers are allocated once for the worst case (every id
distinct) and the de-dup count stays on the device, so the whole step is free of host synchronisation and is captured in a
CUDA graph (``_capture``). Two things ride in that graph besides the step itself: the optimiser chain of the *previous*
step (``attach_optimizer``: clip + noise + cross-client mean + Adam on a side branch beside the frozen trunk), and the
step's host I/O (``enable_host_io``: ids pulled from a host-mapped pinned ring by the first kernel, the loss pushed into
another by the last; ``run_host`` / ``loss_of``).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

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
        """(Re)fill the packed copies **in place**: captured CUDA graphs and in-flight steps hold these pointers, so a
        weight change (load_state_dict, full-state FedAvg) must never move them."""
        b = self.bert
        first = not hasattr(self, "layers")

        def put(store, key, value_fp32, bf16):
            value_fp32 = value_fp32.contiguous()
            if first:
                store[key] = D.cast_bf16(value_fp32) if bf16 else value_fp32.float().clone()
            elif bf16:
                D.cast_bf16(value_fp32, out=store[key])
            else:
                store[key].copy_(value_fp32)

        if first:
            self.emb, self.layers = {}, [dict() for _ in b.transformer.layer]
        put(self.emb, "word", b.embeddings.word_embeddings.weight.data, True)
        put(self.emb, "pos", b.embeddings.position_embeddings.weight.data, True)
        put(self.emb, "g", b.embeddings.LayerNorm.weight.data, False)
        put(self.emb, "b", b.embeddings.LayerNorm.bias.data, False)
        self.word, self.pos, self.emb_g, self.emb_b = self.emb["word"], self.emb["pos"], self.emb["g"], self.emb["b"]
        for blk, L in zip(b.transformer.layer, self.layers):
            a = blk.attention
            put(L, "wqkv", torch.cat([a.q_lin.weight.data, a.k_lin.weight.data, a.v_lin.weight.data], 0), True)
            put(L, "bqkv", torch.cat([a.q_lin.bias.data, a.k_lin.bias.data, a.v_lin.bias.data]), False)
            put(L, "wo", a.out_lin.weight.data, True)
            put(L, "bo", a.out_lin.bias.data, False)
            put(L, "ln1_g", blk.sa_layer_norm.weight.data, False)
            put(L, "ln1_b", blk.sa_layer_norm.bias.data, False)
            put(L, "w1", blk.ffn.lin1.weight.data, True)
            put(L, "b1", blk.ffn.lin1.bias.data, False)
            put(L, "w2", blk.ffn.lin2.weight.data, True)
            put(L, "b2", blk.ffn.lin2.bias.data, False)
            put(L, "ln2_g", blk.output_layer_norm.weight.data, False)
            put(L, "ln2_b", blk.output_layer_norm.bias.data, False)
        # LayerNorm folding (cfg.ln_fold): LN(x)·W^T + b = rstd*(x·(W∘gamma)^T) - rstd*mu*c + (b + W·beta), c = rowsum(W∘gamma)
        # taken from the bf16-rounded product, so a constant row cancels exactly. FFN1 consumes sa_layer_norm of its own
        # block; the QKV projection of block l >= 1 consumes output_layer_norm of block l-1.
        prev = None
        for blk, L in zip(b.transformer.layer, self.layers):
            def fold(key_w, key_c, key_b, w, bias, ln):
                wf, c, bf = ops.reference.fold_layernorm(w, bias, ln.weight.data, ln.bias.data)
                put(L, key_w, wf.float(), True)
                put(L, key_c, c, False)
                put(L, key_b, bf, False)
            fold("w1f", "c1", "b1f", blk.ffn.lin1.weight.data, blk.ffn.lin1.bias.data, blk.sa_layer_norm)
            if prev is not None:
                a = blk.attention
                fold("wqkvf", "cqkv", "bqkvf", torch.cat([a.q_lin.weight.data, a.k_lin.weight.data, a.v_lin.weight.data], 0),
                     torch.cat([a.q_lin.bias.data, a.k_lin.bias.data, a.v_lin.bias.data]), prev.output_layer_norm)
            prev = blk
        self.heads = b.config.n_heads
        self.dim = b.config.dim
        self.ffn = b.config.hidden_dim


def fused_unsupported_reason(model, cfg: FedRecConfig) -> Optional[str]:
    """Why the sm_100a pipeline cannot run this model/config (None = it can). The kernels are built for the reference's
    architecture family: DistilBERT width 768 with 64-wide heads, short titles and histories."""
    bert = model.text_encoder.DistillBert.config
    mha = model.user_encoder.multihead_attention
    q = model.user_encoder.additive_attention.att_fc1.out_features
    T = model.news_tokens.shape[1] if model.news_tokens is not None else cfg.title_len
    checks = [
        (bert.dim == 768 and bert.dim // bert.n_heads == 64, f"DistilBERT width {bert.dim} / head dim {bert.dim // bert.n_heads} (kernels: 768 / 64)"),
        (bert.hidden_dim % 8 == 0, f"FFN width {bert.hidden_dim} is not a multiple of 8"),
        (T <= 64, f"title length {T} > 64 tokens"),
        (cfg.max_his_len <= 64, f"history length {cfg.max_his_len} > 64"),
        (mha.d_k in (16, 20, 32) and mha.d_k == mha.d_v, f"user-encoder head dim {mha.d_k} (kernels: 16, 20, 32)"),
        (mha.n_heads * mha.d_k <= 512 and (mha.n_heads * mha.d_k) % 8 == 0 and mha.d_model % 8 == 0, "user-encoder width"),
        (q <= 256 and q % 8 == 0, f"additive-attention query dim {q}"),
        (cfg.num_candidates <= 32, f"{cfg.num_candidates} candidates per impression > 32"),
    ]
    for ok, why in checks:
        if not ok:
            return why
    return None


class FusedStep:
    def __init__(self, model, cfg: FedRecConfig):
        ops.require_native()
        self.model, self.cfg = model, cfg
        self.dev = model.device
        assert self.dev.type == "cuda"
        self.trunk = PackedTrunk(model.text_encoder.DistillBert)
        self._bufs: Dict[int, Dict[str, torch.Tensor]] = {}
        self._ubufs: Dict[tuple, Dict[str, torch.Tensor]] = {}
        self._step_no = 0
        self._pk: Optional[Dict[str, torch.Tensor]] = None
        self._cache: Optional[Dict[str, torch.Tensor]] = None
        self._side = torch.cuda.Stream(device=self.dev)       # branch for work off the critical path (also inside graphs)
        self._seed_dev = torch.zeros(1, dtype=torch.int64, device=self.dev)   # RNG salt, bumped on the stream every step
        self._graphs: Dict[tuple, dict] = {}
        self._dedup_ws: Optional[torch.Tensor] = None
        self._gen = torch.Generator(device=self.dev)
        self._gen.manual_seed(cfg.seed + 17)
        # optimiser pipelined into the step (attach_optimizer): the chain that consumes step i's gradients runs at the
        # start of step i+1 on its own high-priority branch, beside the frozen trunk
        self._opt = None
        # Normal priority on purpose: with a high-priority branch the captured graph (PDL edges between the trunk kernels,
        # TMEM allocated before griddepcontrol.wait) deadlocked on a B200 (an attention CTA never got its TMEM); the same
        # branch was fine with priority 0, with FEDREC_PDL=0, and without graph capture (profiles/pipeline_ab.md).
        self._opt_stream = torch.cuda.Stream(device=self.dev, priority=int(os.environ.get("FEDREC_OPT_PRIORITY", "0")))
        self._pending = False
        self._opt_forked = False
        self._side_forked = False
        self._io = None                 # host-mapped rings for run_host / loss_of (enable_host_io)
        self._ts = None                 # FEDREC_STEP_TIMESTAMPS=1: %globaltimer at the start and end of every step
        if os.environ.get("FEDREC_STEP_TIMESTAMPS") == "1":
            self._ts = (torch.zeros(8192, dtype=torch.int64, device=self.dev), torch.zeros(1, dtype=torch.int64, device=self.dev))

    # ---- optimiser pipelined into the step -----------------------------------------------------------------
    def attach_optimizer(self, grad_agg, lr: float, b1: float, b2: float, eps: float) -> None:
        """From now on ``run(train=True)`` also *applies* the update: clip + noise + cross-client mean + Adam of step i
        run at the start of step i+1 (or in ``flush()``), on a side branch next to the frozen trunk, which reads
        nothing they write. The arithmetic and its order are those of "step, aggregate, Adam" run back to back
        (upstream: model.update() after the batch loop, main.py:90; DDP overlapping its buckets, main.py:69-70).
        ``grad_agg``: a ``GradAggregator`` with a device chain (transport nvlink or local)."""
        assert grad_agg.on_device_chain, "the pipelined optimiser needs the nvlink or local transport"
        self.flush()
        self._opt = (grad_agg, float(lr), float(b1), float(b2), float(eps))
        grad_agg.symm.set_step(self.model.flat.step)

    def detach_optimizer(self) -> None:
        self.flush()
        self._opt = None

    def sync_step(self) -> None:
        """The host-side Adam step count changed behind the pipeline's back (a snapshot was loaded): copy it into the
        device-side counter the chain's bias corrections use."""
        if self._opt is not None:
            self.flush()
            self._opt[0].symm.set_step(self.model.flat.step)

    def _opt_chain(self, count: bool) -> None:
        agg, lr, b1, b2, eps = self._opt
        if count:
            agg.chain_(lr, b1, b2, eps, gated=True)
        else:                                   # warm-up / capture: launches only, host counters untouched
            step, rnd = agg.flat.step, agg.round
            agg.chain_(lr, b1, b2, eps, gated=True)
            agg.flat.step, agg.round = step, rnd

    def flush(self) -> None:
        """Apply the update of the last step now (before anything reads the trainable weights or the run ends)."""
        if self._opt is not None and self._pending:
            self._opt_chain(count=True)
            self._pending = False

    def repack_trunk(self) -> None:
        self.trunk.repack()
        if self._cache is not None:
            self._cache["valid"].zero_()          # new trunk weights: every cached hidden state is stale

    def _trunk_cache(self) -> Dict[str, torch.Tensor]:
        """[N*T, 768] bf16 table of trunk outputs + per-item valid flags (opt-in, ``cfg.trunk_cache``)."""
        if self._cache is None:
            N, T = self.model.news_tokens.shape
            self._cache = {"hidden": torch.empty(N * T, self.trunk.dim, device=self.dev, dtype=torch.bfloat16),
                           "valid": torch.zeros(N, device=self.dev, dtype=torch.int32)}
        return self._cache

    # ---- bf16 / transposed copies of the *trainable* weights (they change every step) -------------------------
    def _packed(self) -> Dict[str, torch.Tensor]:
        if self._pk is None:
            m = self.model
            te, ue = m.text_encoder, m.user_encoder
            mha, aa = ue.multihead_attention, ue.additive_attention
            HD, Dm, Q = mha.n_heads * mha.d_k, mha.d_model, aa.att_fc1.out_features
            hq, dim, nd = te.additive_attention.att_fc1.out_features, self.trunk.dim, self.cfg.news_dim
            z = lambda *sh, dt=torch.bfloat16: torch.zeros(*sh, device=self.dev, dtype=dt)  # noqa: E731
            self._pk = {"t_w1": z(hq, dim), "t_wfc": z(nd, dim), "t_wfcT": z(dim, nd),
                        "u_wqkv": z(3 * HD, Dm), "u_bqkv": z(3 * HD, dt=torch.float32), "u_wqkvT": z(Dm, 3 * HD),
                        "u_w1": z(Q, HD), "u_w1T": z(HD, Q)}
        return self._pk

    def _pack_weights(self, train: bool) -> None:
        """Casts / transposes of the 16 trainable tensors. Independent of the frozen trunk, so the step runs this on a
        side stream while the trunk computes."""
        m, pk = self.model, self._packed()
        te, ue = m.text_encoder, m.user_encoder
        mha, aa = ue.multihead_attention, ue.additive_attention
        HD = mha.n_heads * mha.d_k
        D.cast_bf16(te.additive_attention.att_fc1.weight.data, out=pk["t_w1"])
        D.cast_bf16(te.fc.weight.data, out=pk["t_wfc"])
        for i, lin in enumerate((mha.W_Q, mha.W_K, mha.W_V)):
            D.cast_bf16(lin.weight.data, out=pk["u_wqkv"][i * HD:(i + 1) * HD])
            pk["u_bqkv"][i * HD:(i + 1) * HD].copy_(lin.bias.data)
        D.cast_bf16(aa.att_fc1.weight.data, out=pk["u_w1"])
        if train:
            D.transpose_cast(te.fc.weight.data, out=pk["t_wfcT"])
            D.transpose_cast(aa.att_fc1.weight.data, out=pk["u_w1T"])
            for i, lin in enumerate((mha.W_Q, mha.W_K, mha.W_V)):
                D.transpose_cast(lin.weight.data, out=pk["u_wqkvT"][:, i * HD:(i + 1) * HD])

    def _fork(self):
        """Start a side branch that sees everything queued on the current stream so far."""
        self._side.wait_stream(torch.cuda.current_stream(self.dev))
        self._side_forked = True
        return torch.cuda.stream(self._side)

    def _join(self) -> None:
        cur = torch.cuda.current_stream(self.dev)
        if self._side_forked:          # only branches that were forked from this stream (a capture admits no others)
            cur.wait_stream(self._side)
            self._side_forked = False
        if self._opt_forked:
            cur.wait_stream(self._opt_stream)
            self._opt_forked = False

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
             "y2": z(R, dim) if self.cfg.ln_fold else None,
             "stats": z(2 * len(self.trunk.layers), R, dim // 64, 2, dt=torch.float32) if self.cfg.ln_fold else None,
             "e": z(R, hq), "pooled": z(K, dim), "alpha": z(R, dt=torch.float32),
             "news_vecs": z(K, self.cfg.news_dim, dt=torch.float32),
             # backward
             "d_news": z(K, self.cfg.news_dim, dt=torch.float32), "d_news_bf": z(K, self.cfg.news_dim),
             "d_pooled": z(K, dim), "d_pre": z(R, hq), "da": z(R, dt=torch.float32),
             # trunk cache bookkeeping
             "miss": z(K, dt=torch.int32), "miss_count": z(1, dt=torch.int32),
             "full_count": torch.full((1,), K, device=dev, dtype=torch.int32)}
        self._bufs[K] = b
        return b

    # ---- trunk + head forward ------------------------------------------------------------------------
    @torch.no_grad()
    def encode(self, uniq: torch.Tensor, count: Optional[torch.Tensor], K: int) -> Dict[str, torch.Tensor]:
        """Inference entry: pending optimiser work is applied first, then ``_encode``."""
        self.flush()
        return self._encode(uniq, count, K)

    @torch.no_grad()
    def _encode(self, uniq: torch.Tensor, count: Optional[torch.Tensor], K: int, packed: bool = False,
                precomputed: bool = False) -> Dict[str, torch.Tensor]:
        """``uniq`` int32 [K] news rows (first ``*count`` valid) → buffers with ``hidden`` (=x) and ``news_vecs``.

        ``packed=True`` means the caller already refreshed the bf16 weight copies on the side stream (joined here,
        after the trunk)."""
        m, tr, b = self.model, self.trunk, self._buffers(K)
        if not packed:
            with self._fork():
                self._pack_weights(train=False)
        if precomputed:
            # cfg.precomputed_news: pooled 768-d features come from the cached table; only the 768->400 projection runs
            if count is None:
                count = b["full_count"]
            D.rows_copy(m.news_features_bf16(), uniq, b["pooled"], None, count, K, 1)
            self._join()
            D.linear_bf16(b["pooled"], self._packed()["t_wfc"], m.text_encoder.fc.bias.data, "bias", out=b["news_vecs"],
                          m_dev=count, m_scale=1)
            return b
        T = m.news_tokens.shape[1]
        x = b["x"]
        want_uniq, want_count = uniq, count
        cache = self._trunk_cache() if self.cfg.trunk_cache else None
        if cache is not None:
            # only news items never seen before go through the trunk; everything else is gathered from HBM below
            if count is None:
                count = b["full_count"]
            D.cache_select(uniq, count, K, cache["valid"], b["miss"], b["miss_count"])
            uniq, count = b["miss"], b["miss_count"]
        D.bert_embed_ln(m.news_tokens, uniq, count, K, T, tr.word, tr.pos, tr.emb_g, tr.emb_b, x)
        if self.cfg.ln_fold:
            self._trunk_folded(b, uniq, count, K, T)
        else:
            self._trunk_plain(b, uniq, count, K, T)
        if cache is not None:
            D.rows_copy(x, None, cache["hidden"], uniq, count, K, T, valid=cache["valid"])      # new items → cache
            uniq, count = want_uniq, (want_count if want_count is not None else b["full_count"])
            D.rows_copy(cache["hidden"], uniq, x, None, count, K, T)                            # this step's items ← cache
        te = m.text_encoder
        aa = te.additive_attention
        pk = self._packed()
        self._join()                                   # packed weights are ready
        D.linear_bf16(x, pk["t_w1"], aa.att_fc1.bias.data, "tanh", out=b["e"], m_dev=count, m_scale=T)
        mask = m.news_mask if te.mask_padding else None
        D.text_pool(x, b["e"], aa.att_fc2.weight.data.view(-1), aa.att_fc2.bias.data, mask, uniq, count, K, T,
                    b["pooled"], b["alpha"])
        D.linear_bf16(b["pooled"], pk["t_wfc"], te.fc.bias.data, "bias", out=b["news_vecs"], m_dev=count, m_scale=1)
        return b

    def _trunk_plain(self, b, uniq, count, K: int, T: int) -> None:
        """The six blocks with LayerNorm as a kernel of its own (``--ln_fold false``); ``x`` holds the block input / output."""
        m, tr = self.model, self.trunk
        x, y, ctx, qkv, ffn = b["x"], b["y"], b["ctx"], b["qkv"], b["ffn"]
        for L in tr.layers:
            D.linear_bf16(x, L["wqkv"], L["bqkv"], "bias", out=qkv, m_dev=count, m_scale=T)
            D.bert_attention(qkv, m.news_mask, uniq, count, K, T, tr.heads, ctx)
            D.linear_bf16(ctx, L["wo"], L["bo"], "resid", resid=x, out=y, m_dev=count, m_scale=T)
            D.layernorm_bf16(y, L["ln1_g"], L["ln1_b"], out=x, count=count, rows_per_item=T)
            D.linear_bf16(x, L["w1"], L["b1"], "gelu", out=ffn, m_dev=count, m_scale=T)
            D.linear_bf16(ffn, L["w2"], L["b2"], "resid", resid=x, out=y, m_dev=count, m_scale=T)
            D.layernorm_bf16(y, L["ln2_g"], L["ln2_b"], out=x, count=count, rows_per_item=T)

    def _trunk_folded(self, b, uniq, count, K: int, T: int) -> None:
        """The six blocks with every LayerNorm but the last folded into the GEMMs around it (GEMM notes).

        ``y`` / ``y2`` hold the *un-normalised* block outputs; their row statistics travel in ``stats`` (one (sum, sum of
        squares) pair per 64-column slice, stored by the residual epilogues without atomics, so the step stays
        bit-reproducible) and each consumer applies the normalisation in its own epilogue."""
        m, tr = self.model, self.trunk
        x, y1, y2, ctx, qkv, ffn, st = b["x"], b["y"], b["y2"], b["ctx"], b["qkv"], b["ffn"], b["stats"]
        kw = dict(m_dev=count, m_scale=T)
        n = len(tr.layers)
        for l, L in enumerate(tr.layers):
            st1, st2 = st[2 * l], st[2 * l + 1]
            if l == 0:
                D.linear_bf16(x, L["wqkv"], L["bqkv"], "bias", out=qkv, **kw)
                D.bert_attention(qkv, m.news_mask, uniq, count, K, T, tr.heads, ctx)
                D.linear_bf16_fold(ctx, L["wo"], L["bo"], "resid", y1, resid=x, stats_out=st1, **kw)
            else:
                P, pst = tr.layers[l - 1], st[2 * l - 1]
                D.linear_bf16_fold(y2, L["wqkvf"], L["bqkvf"], "bias", qkv, a_stats=pst, col_c=L["cqkv"], **kw)
                D.bert_attention(qkv, m.news_mask, uniq, count, K, T, tr.heads, ctx)
                D.linear_bf16_fold(ctx, L["wo"], L["bo"], "resid", y1, resid=y2, r_stats=pst, r_gamma=P["ln2_g"],
                                   r_beta=P["ln2_b"], stats_out=st1, **kw)
            D.linear_bf16_fold(y1, L["w1f"], L["b1f"], "gelu", ffn, a_stats=st1, col_c=L["c1"], **kw)
            D.linear_bf16_fold(ffn, L["w2"], L["b2"], "resid", y2, resid=y1, r_stats=st1, r_gamma=L["ln1_g"],
                               r_beta=L["ln1_b"], stats_out=st2 if l + 1 < n else None, **kw)
        last = tr.layers[-1]
        D.layernorm_bf16(y2, last["ln2_g"], last["ln2_b"], out=x, count=count, rows_per_item=T)

    def _ids(self, cand, his, ids: Optional[torch.Tensor] = None):
        if ids is None:
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

    # ---- user encoder + loss, forward (and backward when training) -----------------------------------------
    def _user_buffers(self, B: int, L: int, C: int) -> Dict[str, torch.Tensor]:
        key = (B, L, C)
        if key in self._ubufs:
            return self._ubufs[key]
        m = self.model
        mha = m.user_encoder.multihead_attention
        D, H, dk = mha.d_model, mha.n_heads, mha.d_k
        Q = m.user_encoder.additive_attention.att_fc1.out_features
        bf, dev = torch.bfloat16, self.dev
        z = lambda *s, dt=bf: torch.zeros(*s, device=dev, dtype=dt)  # noqa: E731
        R = B * L
        u = {"X": z(R, D), "qkv": z(R, 3 * H * dk), "ctx": z(R, H * dk), "stats": z(B * H * L, 2, dt=torch.float32),
             "E": z(R, Q), "score": z(B, C, dt=torch.float32), "loss": z((), dt=torch.float32),
             "user": z(B, H * dk, dt=torch.float32), "d_ctx_pool": z(R, H * dk), "d_pre": z(R, Q), "d_ctx": z(R, H * dk),
             "dqkv": z(R, 3 * H * dk), "dX": z(R, D, dt=torch.float32)}
        self._ubufs[key] = u
        return u

    def _user_forward(self, b, inv, his, B, L, C, train: bool, seed: int):
        m, cfg = self.model, self.cfg
        ue = m.user_encoder
        mha, aa = ue.multihead_attention, ue.additive_attention
        H, dk = mha.n_heads, mha.d_k
        u = self._user_buffers(B, L, C)
        nv = b["news_vecs"]
        inv_his, inv_cand = inv[: B * L], inv[B * L:]
        p = float(ue.dropout_rate) if train else 0.0
        use_mask = bool(cfg.mask_padding)
        his_flat = his.reshape(-1).to(torch.int32) if use_mask else None
        pk = self._packed()
        D.ue_gather_dropout(nv, inv_his, u["X"], p, seed, self._seed_dev)
        D.linear_bf16(u["X"], pk["u_wqkv"], pk["u_bqkv"], "bias", out=u["qkv"])
        D.ue_mha_fwd(u["qkv"], his_flat, u["ctx"], u["stats"], B, L, H, dk, use_mask)
        D.linear_bf16(u["ctx"], pk["u_w1"], aa.att_fc1.bias.data, "tanh", out=u["E"])
        fl = m.flat
        if train:
            D.ue_pool_score(u["ctx"], u["E"], aa.att_fc2.weight.data.view(-1), aa.att_fc2.bias.data, nv, inv_cand,
                            his_flat, u["score"], u["loss"], u["user"], u["d_ctx_pool"], u["d_pre"],
                            fl.view("user_encoder.additive_attention.att_fc2.weight", grad=True).view(-1),
                            fl.view("user_encoder.additive_attention.att_fc2.bias", grad=True), b["d_news"], B, L, C,
                            use_mask, True, cfg.news_grad_noise, seed + 1, self._seed_dev)
        else:
            D.ue_pool_score(u["ctx"], u["E"], aa.att_fc2.weight.data.view(-1), aa.att_fc2.bias.data, nv, inv_cand,
                            his_flat, u["score"], u["loss"], u["user"], None, None, None, None, None, B, L, C,
                            use_mask, False)
        return u, p, his_flat, use_mask

    def _user_backward(self, u, b, inv, his_flat, use_mask, B, L, C, p, seed):
        """Critical path: data gradients down to d_news. Weight / bias gradients go to the side branch."""
        m, cfg = self.model, self.cfg
        ue = m.user_encoder
        mha = ue.multihead_attention
        H, dk = mha.n_heads, mha.d_k
        HD = H * dk
        fl, pk = m.flat, self._packed()
        g = lambda name: fl.view(name, grad=True)  # noqa: E731
        inv_his = inv[: B * L]
        with self._fork():                                   # needs d_pre, ctx
            D.gemm_grad(u["d_pre"], u["ctx"], g("user_encoder.additive_attention.att_fc1.weight"))
            D.colsum(u["d_pre"], g("user_encoder.additive_attention.att_fc1.bias"))
        # additive attention fc1 data gradient (+ pooling path as the residual), then attention
        D.linear_bf16(u["d_pre"], pk["u_w1T"], None, "resid", resid=u["d_ctx_pool"], out=u["d_ctx"])
        D.ue_mha_bwd(u["qkv"], u["d_ctx"], u["stats"], his_flat, u["dqkv"], B, L, H, dk, use_mask)
        with self._fork():                                   # needs dqkv, X
            for i, nm in enumerate(("W_Q", "W_K", "W_V")):
                sl = u["dqkv"][:, i * HD:(i + 1) * HD]
                D.gemm_grad(sl, u["X"], g(f"user_encoder.multihead_attention.{nm}.weight"))
                D.colsum(sl, g(f"user_encoder.multihead_attention.{nm}.bias"))
        D.linear_bf16(u["dqkv"], pk["u_wqkvT"], None, "bias", out=u["dX"])
        D.ue_scatter_dx(u["dX"], inv_his, b["d_news"], p, seed, cfg.news_grad_noise, seed + 2, self._seed_dev)

    def _text_head_backward(self, b, count, K):
        """d_news [K, D] → gradients of text_encoder.fc and text_encoder.additive_attention (trunk frozen)."""
        m = self.model
        te = m.text_encoder
        aa = te.additive_attention
        T = m.news_tokens.shape[1]
        fl, pk = m.flat, self._packed()
        g = lambda name: fl.view(name, grad=True)  # noqa: E731
        D.cast_bf16(b["d_news"], out=b["d_news_bf"])
        with self._fork():                                   # needs d_news(_bf), pooled
            D.gemm_grad(b["d_news_bf"], b["pooled"], g("text_encoder.fc.weight"), k_dev=count, k_scale=1)
            D.colsum(b["d_news"], g("text_encoder.fc.bias"), m_dev=count, m_scale=1)
        if self.cfg.precomputed_news:                        # the additive attention is frozen into the feature cache
            self._join()
            return
        D.linear_bf16(b["d_news_bf"], pk["t_wfcT"], None, "bias", out=b["d_pooled"], m_dev=count, m_scale=1)
        D.text_pool_bwd(b["x"], b["e"], b["alpha"], b["d_pooled"], aa.att_fc2.weight.data.view(-1), count, K, T,
                        b["d_pre"], b["da"], None, g("text_encoder.additive_attention.att_fc2.bias"), None)
        with self._fork():                                   # needs da, d_pre: the two reductions over all rows
            D.colsum_weighted(b["e"], b["da"], g("text_encoder.additive_attention.att_fc2.weight").view(-1),
                              m_dev=count, m_scale=T)
            D.colsum(b["d_pre"], g("text_encoder.additive_attention.att_fc1.bias"), m_dev=count, m_scale=T)
        D.gemm_grad(b["d_pre"], b["x"], g("text_encoder.additive_attention.att_fc1.weight"), k_dev=count, k_scale=T)
        self._join()                                         # every gradient is in the flat buffer

    # ---- host I/O through host-mapped rings --------------------------------------------------------------------
    def enable_host_io(self, n_ids: int, slots: int = 4) -> None:
        """Rings in pinned, device-visible host memory: ``slots`` x ``n_ids`` int32 going in, ``slots`` x {loss, sequence
        number} coming out. The captured step starts with a kernel that pulls its slot over PCIe and ends with one that
        pushes the loss back, so a step costs the host one memcpy into pinned memory, one graph launch and a poll of
        host memory: no copy-engine hop, no event, no stream synchronisation."""
        import ctypes
        import numpy as np
        lib = ops.require_native()
        if self._io is not None and self._io["n"] == n_ids and self._io["slots"] == slots:
            return
        assert self._io is None, "host I/O rings are sized once"
        hp, dp = ctypes.c_void_p(), ctypes.c_void_p()
        ho, do = ctypes.c_void_p(), ctypes.c_void_p()
        n_pad = (n_ids + 3) // 4 * 4                     # slots are read with 16-byte loads
        with torch.cuda.device(self.dev):
            D.check(lib.fr_host_mapped_alloc(slots * n_pad * 4, ctypes.byref(hp), ctypes.byref(dp)), "fr_host_mapped_alloc")
            D.check(lib.fr_host_mapped_alloc(slots * 16, ctypes.byref(ho), ctypes.byref(do)), "fr_host_mapped_alloc")
        ring = np.ctypeslib.as_array((ctypes.c_int32 * (slots * n_pad)).from_address(hp.value)).reshape(slots, n_pad)
        out = np.ctypeslib.as_array((ctypes.c_uint32 * (slots * 4)).from_address(ho.value)).reshape(slots, 4)
        self._io = {"n": n_ids, "n_pad": n_pad, "slots": slots, "ring": ring, "ring_dev": dp.value, "out": out, "out_dev": do.value,
                    "ctr": torch.zeros(1, dtype=torch.int64, device=self.dev), "issued": 0, "confirmed": 0}

    def run_host(self, ids, B: int, C: int, L: int) -> int:
        """One training step whose ids come from host memory (``ids``: int32 ``[B*L + B*C]``, histories first; numpy or
        CPU tensor). Returns a ticket for ``loss_of``. Call ``enable_host_io`` first. The host may run at most
        ``slots - 1`` steps ahead of the results it has read; beyond that this call waits for the oldest one."""
        import numpy as np
        io = self._io
        assert io is not None and io["n"] == B * (L + C), "enable_host_io(B * (L + C)) first"
        if io["issued"] - io["confirmed"] >= io["slots"] - 1:
            self.loss_of(io["confirmed"] + 1)
        src = ids.numpy() if isinstance(ids, torch.Tensor) else np.asarray(ids)
        np.copyto(io["ring"][io["issued"] % io["slots"], : io["n"]], src.reshape(-1), casting="no")
        io["issued"] += 1
        if self.cfg.cuda_graph:
            self._replay(self._graph_for(B, C, L, hostio=True))
        else:
            self._step_eager(None, None, True, hostio=(B, C, L))
        return io["issued"]

    def loss_of(self, ticket: int, timeout_s: float = 120.0) -> float:
        """Loss of the step ``run_host`` returned ``ticket`` for; waits (polling host memory) until it has landed."""
        import time
        io = self._io
        assert 0 < ticket <= io["issued"] and ticket > io["issued"] - io["slots"], "result already overwritten"
        slot = io["out"][(ticket - 1) % io["slots"]]
        t0 = None
        while int(slot[1]) != (ticket & 0xFFFFFFFF):
            if t0 is None:
                t0 = time.monotonic()
            elif time.monotonic() - t0 > timeout_s:
                torch.cuda.synchronize(self.dev)           # surfaces a device-side failure as an exception
                raise TimeoutError(f"step {ticket}: no result after {timeout_s:.0f}s")
        io["confirmed"] = max(io["confirmed"], ticket)
        return float(slot[:1].view("float32")[0])

    # ---- full step -----------------------------------------------------------------------------------------
    def _graph_key(self, B: int, C: int, L: int, hostio: bool = False) -> tuple:
        cfg = self.cfg      # everything a captured launch bakes in as a constant is part of the key
        opt = None
        if self._opt is not None:
            agg = self._opt[0]
            opt = (id(agg), agg.clip, agg.noise_std, len(agg.participants), bool(agg.symm.mc_ptr)) + self._opt[1:]
        return (B, C, L, float(cfg.news_grad_noise), bool(cfg.mask_padding), bool(cfg.dedup), bool(cfg.trunk_cache),
                bool(cfg.ln_fold), float(self.model.user_encoder.dropout_rate), opt, hostio, bool(cfg.precomputed_news))

    def run(self, cand: torch.Tensor, his: torch.Tensor, train: bool = True) -> torch.Tensor:
        """One step. With ``cfg.cuda_graph`` the whole kernel sequence (≈90 launches) is captured once per batch
        shape and replayed; the ids go into the graph's one static input buffer (histories first, then candidates)
        and a device-side counter salts the dropout / noise streams so every replay draws fresh randomness.
        With an attached optimiser (``attach_optimizer``) the step also applies the previous step's update."""
        if not (train and self.cfg.cuda_graph):
            return self._step_eager(cand, his, train)
        B, C = cand.shape
        L = his.shape[1]
        g = self._graph_for(B, C, L)
        if cand.dtype == torch.int32 and his.dtype == torch.int32:
            torch.cat((his.reshape(-1), cand.reshape(-1)), out=g["ids"])        # one kernel, no copy-engine hop
        else:
            g["ids"][: B * L].copy_(his.reshape(-1), non_blocking=True)
            g["ids"][B * L:].copy_(cand.reshape(-1), non_blocking=True)
        return self._replay(g)

    def run_packed(self, ids: torch.Tensor, B: int, C: int, L: int) -> torch.Tensor:
        """``run`` for callers that already hold the step's ids in the graph's input layout: int32 ``[B*L + B*C]``,
        the B histories first, then the B candidate lists. ``ids`` may be pinned host memory: the step then starts
        with exactly one host-to-device copy straight into the static buffer."""
        assert ids.dtype == torch.int32 and ids.numel() == B * (L + C)
        if not self.cfg.cuda_graph:
            d = ids.to(self.dev, non_blocking=True)
            return self._step_eager(d[B * L:].view(B, C), d[: B * L].view(B, L), True)
        g = self._graph_for(B, C, L)
        g["ids"].copy_(ids, non_blocking=True)
        return self._replay(g)

    def _count_update(self) -> None:
        agg = self._opt[0]
        agg.symm.check()
        agg.round += 1
        agg.flat.step += 1

    def _replay(self, g) -> torch.Tensor:
        pipelined = self._opt is not None
        if pipelined and self._pending:
            self._count_update()                        # this replay applies the previous step's update
        g["graph"].replay()
        D.count_launches(g["launches"])
        if pipelined:
            self._pending = True
        return g["loss"]

    def _step_eager(self, cand, his, train, hostio=None) -> torch.Tensor:
        pipelined = train and self._opt is not None
        if pipelined and self._pending:
            self._count_update()
        if hostio is not None:
            B, C, L = hostio
            ids = self._io.setdefault("ids_eager", torch.zeros(self._io["n_pad"], dtype=torch.int32, device=self.dev))[: B * (L + C)]
            loss = self._run_eager(ids[B * L:].view(B, C), ids[: B * L].view(B, L), train, ids=ids, hostio=True)
        else:
            loss = self._run_eager(cand, his, train)
        if pipelined:
            self._pending = True
        return loss

    def _graph_for(self, B: int, C: int, L: int, hostio: bool = False):
        key = self._graph_key(B, C, L, hostio)
        g = self._graphs.get(key)
        if g is None:
            g = self._capture(B, C, L, key, hostio)
        return g

    def _capture(self, B: int, C: int, L: int, key, hostio: bool = False):
        self.flush()                                        # no pending update may leak into the warm-up runs
        ids = torch.ones((B * (L + C) + 3) // 4 * 4, dtype=torch.int32, device=self.dev)[: B * (L + C)]   # padded storage
        static_h, static_c = ids[: B * L].view(B, L), ids[B * L:].view(B, C)
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):                       # warm-up: buffers, attributes, tensor maps, module loads
            for _ in range(2):
                self._run_eager(static_c, static_h, True, warm=True, ids=ids if hostio else None, hostio=hostio)
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        if hostio:      # the warm-up passes consumed ring slots and published results: put the counter back and forget
            io = self._io   # what they published (results of earlier steps that the caller has not read yet stay)
            io["ctr"].fill_(io["issued"] - 1)
            torch.cuda.synchronize(self.dev)
            io["out"][io["out"][:, 1] >= (io["issued"] & 0xFFFFFFFF), 1] = 0
        graph = torch.cuda.CUDAGraph()
        n0 = D.launch_count()
        with torch.cuda.graph(graph):
            loss = self._run_eager(static_c, static_h, True, ids=ids, hostio=hostio)
        g = {"graph": graph, "ids": ids, "cand": static_c, "his": static_h, "loss": loss,
             "launches": D.launch_count() - n0}
        self._graphs[key] = g
        return g

    @torch.no_grad()
    def _run_eager(self, cand: torch.Tensor, his: torch.Tensor, train: bool = True, warm: bool = False,
                   ids: Optional[torch.Tensor] = None, hostio: bool = False) -> torch.Tensor:
        """The step's launches on the current stream. ``warm``: a warm-up pass before capture — the optimiser chain is
        launched (its kernels return at once: nothing is pending) but gradients are not marked pending.
        ``ids``: the already concatenated int32 ids (histories, then candidates) when the caller has them."""
        m = self.model
        B, C = cand.shape
        L = his.shape[1]
        self._step_no += 1
        seed = self.cfg.seed * 1000003 + 3 * self._step_no
        if self._ts is not None and train:
            D.timestamp(*self._ts)
        D.counter_advance(self._seed_dev, 0x9E3779B97F4A7C15 & (2 ** 63 - 1))
        if hostio:                                           # this step's ids: host ring slot -> static id buffer
            io = self._io
            D.hostio_fetch(ids, io["ring_dev"], io["n_pad"], io["slots"], io["ctr"])     # ids' storage is padded to n_pad
        opt = self._opt if train else None
        cur = torch.cuda.current_stream(self.dev)
        if opt is not None:
            # the previous step's update, then everything that depends on it (zeroed gradients, bf16 weight copies),
            # on a high-priority branch beside the frozen trunk
            self._opt_stream.wait_stream(cur)
            with torch.cuda.stream(self._opt_stream):
                self._opt_chain(count=False)
                m.flat.zero_grad()
                self._pack_weights(train)
            self._opt_forked = True
        else:
            with self._fork():                               # trainable-weight copies: off the critical path
                self._pack_weights(train)
        ids, uniq, count, inv, K = self._ids(cand, his, ids)
        if train and opt is None:
            m.flat.zero_grad()
        pre = bool(self.cfg.precomputed_news)
        if pre and m.news_features is None:
            raise RuntimeError("cfg.precomputed_news: call model.build_news_features() (or set_news_features) first")
        b = self._encode(uniq, count, K, packed=True, precomputed=pre)
        if train:
            b["d_news"].zero_()
        u, p, his_flat, use_mask = self._user_forward(b, inv, his, B, L, C, train, seed)
        if train:
            self._user_backward(u, b, inv, his_flat, use_mask, B, L, C, p, seed)
            self._text_head_backward(b, count if count is not None else None, K)
            if opt is not None:
                opt[0].symm.mark_pending(arm=not warm)       # gradients complete: arm the chain of the next step
        self.last_scores = u["score"]
        if self._ts is not None and train:
            D.timestamp(*self._ts)
        if hostio:
            D.hostio_publish(self._io["out_dev"], u["loss"], self._io["slots"], self._io["ctr"])
        return u["loss"].clone()

    @torch.no_grad()
    def scores(self, cand: torch.Tensor, his: torch.Tensor, chunk: int = 64) -> torch.Tensor:
        """Inference scores ``[B, C]``. Large batches are processed ``chunk`` users at a time so the activation buffers
        (sized for the worst case of every id being distinct) stay at a couple of fixed sizes."""
        B, C = cand.shape
        if B > chunk:
            return torch.cat([self.scores(cand[i:i + chunk], his[i:i + chunk], chunk) for i in range(0, B, chunk)])
        L = his.shape[1]
        self.flush()
        with self._fork():
            self._pack_weights(False)
        ids, uniq, count, inv, K = self._ids(cand, his)
        b = self._encode(uniq, count, K, packed=True,
                         precomputed=bool(self.cfg.precomputed_news) and self.model.news_features is not None)
        u, _, _, _ = self._user_forward(b, inv, his, B, L, C, False, 0)
        return u["score"].clone()
