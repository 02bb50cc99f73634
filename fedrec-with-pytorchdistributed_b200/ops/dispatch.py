"""Device dispatch: CUDA tensors → sm_100a kernels (mandatory), CPU tensors → ``ops.reference``."""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import reference as ref
from .native import check, require_native

_LAUNCHES = 0   # kernels launched through this layer (bench.py reports it as gpu_launches)


def launch_count() -> int:
    return _LAUNCHES


def count_launches(n: int = 1) -> None:
    global _LAUNCHES
    _LAUNCHES += n


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.contiguous().float()
    return t


# ----------------------------------------------------------------------------------------------
class _ScoreCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cand, user, labels):
        lib = require_native()
        cand, user = _f32c(cand), _f32c(user)
        B, C, D = cand.shape
        score = torch.empty(B, C, device=cand.device, dtype=torch.float32)
        loss = torch.empty((), device=cand.device, dtype=torch.float32)
        d_cand = torch.empty_like(cand)
        d_user = torch.empty_like(user)
        lab = labels.to(torch.int32).contiguous() if labels is not None else None
        check(lib.fr_score_ce(cand.data_ptr(), user.data_ptr(), _p(lab), score.data_ptr(), None, loss.data_ptr(),
                              d_cand.data_ptr(), d_user.data_ptr(), 1.0, B, C, D, stream_ptr(cand.device)), "fr_score_ce")
        count_launches(1)
        ctx.save_for_backward(d_cand, d_user)
        ctx.mark_non_differentiable(score)
        return loss, score

    @staticmethod
    def backward(ctx, g_loss, _g_score):
        d_cand, d_user = ctx.saved_tensors
        return d_cand * g_loss, d_user * g_loss, None


def score_ce(cand: torch.Tensor, user: torch.Tensor, labels: Optional[torch.Tensor] = None):
    if cand.is_cuda:
        return _ScoreCE.apply(cand, user, labels)
    return ref.score_ce(cand, user, labels)


def segment_sum(vals: torch.Tensor, seg: torch.Tensor, num_segments: int) -> torch.Tensor:
    if not vals.is_cuda:
        return ref.segment_sum(vals, seg, num_segments)
    lib = require_native()
    vals = _f32c(vals)
    seg = seg.to(torch.int32).contiguous()
    K, D = vals.shape
    out = torch.empty(num_segments, D, device=vals.device, dtype=torch.float32)
    check(lib.fr_segment_sum(vals.data_ptr(), seg.data_ptr(), out.data_ptr(), K, num_segments, D,
                             stream_ptr(vals.device)), "fr_segment_sum")
    count_launches(1)
    return out


def adam_flat(flat, lr: float, b1: float, b2: float, eps: float, grad_scale: float = 1.0) -> None:
    """One Adam step over ``flat.param`` using ``flat.grad`` (torch.optim.Adam defaults)."""
    flat.step += 1
    if not flat.param.is_cuda:
        with torch.no_grad():
            ref.adam_(flat.param, flat.grad, flat.exp_avg, flat.exp_avg_sq, flat.step, lr, b1, b2, eps, grad_scale)
        return
    lib = require_native()
    bc1 = 1.0 - b1 ** flat.step
    bc2 = 1.0 - b2 ** flat.step
    check(lib.fr_adam_flat(flat.param.data_ptr(), flat.grad.data_ptr(), flat.exp_avg.data_ptr(),
                           flat.exp_avg_sq.data_ptr(), flat.numel, lr, b1, b2, eps, bc1, bc2, None, grad_scale,
                           stream_ptr(flat.param.device)), "fr_adam_flat")
    count_launches(1)


def grad_sq_norm(x: torch.Tensor) -> torch.Tensor:
    if not x.is_cuda:
        return (x.float() ** 2).sum()
    lib = require_native()
    out = torch.empty((), device=x.device, dtype=torch.float32)
    check(lib.fr_sq_norm(x.data_ptr(), x.numel(), out.data_ptr(), stream_ptr(x.device)), "fr_sq_norm")
    count_launches(1)
    return out


def clip_noise_(x: torch.Tensor, clip: float, noise_std: float, seed: int = 0, stream_id: int = 0,
                generator=None) -> torch.Tensor:
    """In place: scale ``x`` to L2 norm ≤ ``clip`` (0 = off), then add ``N(0, noise_std²)``."""
    if clip <= 0 and noise_std <= 0:
        return x
    if not x.is_cuda:
        if generator is None and noise_std > 0:      # same contract as the kernel: the noise is a function of (seed, stream)
            generator = torch.Generator().manual_seed((int(seed) * 1000003 + int(stream_id) * 7919 + 12345) & (2 ** 63 - 1))
        return ref.clip_noise_(x, clip, noise_std, generator)
    lib = require_native()
    sq = grad_sq_norm(x) if clip > 0 else None
    check(lib.fr_clip_noise(x.data_ptr(), x.numel(), _p(sq), float(clip), float(noise_std), int(seed) & (2**64 - 1),
                            int(stream_id) & 0xFFFFFFFF, stream_ptr(x.device)), "fr_clip_noise")
    count_launches(1)
    return x


def dedup(ids: torch.Tensor, table_rows: int, workspace: Optional[torch.Tensor] = None
          ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Sync-free unique: returns ``(uniq[K] padded with 0, count[1] int32 device, inv[K])``."""
    flat = ids.reshape(-1)
    K = flat.numel()
    if not ids.is_cuda:
        u, inv = ref.dedup(flat)
        uniq = torch.zeros(K, dtype=torch.int32)
        uniq[: u.numel()] = u.to(torch.int32)
        return uniq, torch.tensor([u.numel()], dtype=torch.int32), inv.to(torch.int32)
    lib = require_native()
    flat = flat.to(torch.int32).contiguous()
    need = lib.fr_dedup_workspace_ints(table_rows)
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty(need, dtype=torch.int32, device=ids.device)
    uniq = torch.empty(K, dtype=torch.int32, device=ids.device)
    inv = torch.empty(K, dtype=torch.int32, device=ids.device)
    count = torch.empty(1, dtype=torch.int32, device=ids.device)
    check(lib.fr_dedup(flat.data_ptr(), K, table_rows, workspace.data_ptr(), uniq.data_ptr(), count.data_ptr(),
                       inv.data_ptr(), stream_ptr(ids.device)), "fr_dedup")
    count_launches(5)
    return uniq, count, inv


_EPI = {"none": 0, "bias": 0, "gelu": 1, "resid": 2, "tanh": 3}


def linear_bf16(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, epi: str = "bias",
                resid: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                out_dtype: torch.dtype = torch.bfloat16, block_n: int = 0,
                m_dev: Optional[torch.Tensor] = None, m_scale: int = 1, cta_group: int = 0,
                tma_epilogue: int = 0) -> torch.Tensor:
    """``epi(x @ w.T + bias)`` on the tcgen05 GEMM. ``x [M,K]`` and ``w [N,K]`` bf16 with unit inner stride.

    ``cta_group`` (0 auto / 1 / 2) and ``tma_epilogue`` (0 auto / -1 direct / 1 staged TMA) pin the kernel variant."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    assert x.dim() == 2 and w.dim() == 2 and x.stride(1) == 1 and w.stride(1) == 1
    lib = require_native()
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, device=x.device, dtype=out_dtype)
    assert out.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.is_contiguous()
    if resid is not None:
        assert resid.dtype == torch.bfloat16 and resid.stride(1) == 1
    check(lib.fr_gemm_bf16_ex(x.data_ptr(), w.data_ptr(), _p(bias), _p(resid), out.data_ptr(),
                              1 if out.dtype == torch.float32 else 0, M, N, K, x.stride(0), w.stride(0), out.stride(0),
                              resid.stride(0) if resid is not None else 0, _EPI[epi], block_n, _p(m_dev), m_scale,
                              cta_group, tma_epilogue, stream_ptr(x.device)), "fr_gemm_bf16")
    count_launches(1)
    return out


def linear_bf16_fold(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], epi: str, out: torch.Tensor,
                     resid: Optional[torch.Tensor] = None, m_dev: Optional[torch.Tensor] = None, m_scale: int = 1,
                     a_stats: Optional[torch.Tensor] = None, col_c: Optional[torch.Tensor] = None,
                     r_stats: Optional[torch.Tensor] = None, r_gamma: Optional[torch.Tensor] = None,
                     r_beta: Optional[torch.Tensor] = None, stats_out: Optional[torch.Tensor] = None,
                     ln_eps: float = 1e-12) -> torch.Tensor:
    """``linear_bf16`` with LayerNorm folded around it (see ``fr_gemm_bf16_fold`` in the header).

    ``a_stats``/``col_c``: the input is an un-normalised tensor whose LayerNorm is applied algebraically (``w`` must
    already be scaled by gamma). ``r_stats``/``r_gamma``/``r_beta``: the residual is normalised on the fly.
    Statistics tensors are ``[M, width // 64, 2]`` fp32: a (sum, sum of squares) pair per 64-column slice of the row they
    describe; ``stats_out`` is written (not accumulated), so it needs no zeroing."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and out.dtype == torch.bfloat16
    assert x.stride(1) == 1 and w.stride(1) == 1 and out.stride(1) == 1
    for t in (a_stats, r_stats, stats_out):
        assert t is None or (t.dtype == torch.float32 and t.is_contiguous() and t.shape[-1] == 2)
    lib = require_native()
    M, K = x.shape
    N = w.shape[0]
    check(lib.fr_gemm_bf16_fold(x.data_ptr(), w.data_ptr(), _p(bias), _p(resid), out.data_ptr(), M, N, K, x.stride(0),
                                w.stride(0), out.stride(0), resid.stride(0) if resid is not None else 0, _EPI[epi],
                                _p(m_dev), m_scale, _p(a_stats), _p(col_c), _p(r_stats), _p(r_gamma), _p(r_beta),
                                _p(stats_out), float(ln_eps), stream_ptr(x.device)), "fr_gemm_bf16_fold")
    count_launches(1)
    return out


# ---- DistilBERT trunk / text head pieces (CUDA only; the CPU path is the nn.Module code) -------------
def bert_embed_ln(tokens, items, count, n_items, T, word, pos, gamma, beta, out, eps=1e-12):
    lib = require_native()
    check(lib.fr_bert_embed_ln(tokens.data_ptr(), _p(items), _p(count), n_items, T, word.data_ptr(), pos.data_ptr(),
                               gamma.data_ptr(), beta.data_ptr(), out.data_ptr(), eps, word.shape[1],
                               stream_ptr(out.device)), "fr_bert_embed_ln")
    count_launches(1)
    return out


def layernorm_bf16(x, gamma, beta, out=None, count=None, rows_per_item=1, eps=1e-12):
    lib = require_native()
    out = torch.empty_like(x) if out is None else out
    check(lib.fr_layernorm_bf16(x.data_ptr(), out.data_ptr(), gamma.data_ptr(), beta.data_ptr(), x.shape[0], _p(count),
                                rows_per_item, eps, x.shape[1], stream_ptr(x.device)), "fr_layernorm_bf16")
    count_launches(1)
    return out


def bert_attention(qkv, mask_table, items, count, n_items, T, heads, out, impl: Optional[str] = None):
    """Title self-attention. ``impl``: "tc" (tcgen05, default) or "simt" (CUDA-core kernel); FEDREC_ATTN overrides."""
    lib = require_native()
    impl = impl or os.environ.get("FEDREC_ATTN", "tc")
    fn = lib.fr_bert_attention_tc if impl == "tc" else lib.fr_bert_attention
    check(fn(qkv.data_ptr(), _p(mask_table), _p(items), _p(count), n_items, T, heads,
                                qkv.shape[1] // (3 * heads), out.data_ptr(), stream_ptr(out.device)), "fr_bert_attention")
    count_launches(1)
    return out


def text_pool(hidden, e, w2, b2, mask_table, items, count, n_items, T, pooled, alpha):
    lib = require_native()
    check(lib.fr_text_pool(hidden.data_ptr(), e.data_ptr(), w2.data_ptr(), b2.data_ptr(), _p(mask_table), _p(items),
                           _p(count), n_items, T, e.shape[1], hidden.shape[1], pooled.data_ptr(), _p(alpha),
                           stream_ptr(hidden.device)), "fr_text_pool")
    count_launches(1)
    return pooled


def cast_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 → bf16 copy with our kernel (weight packing)."""
    lib = require_native()
    x = x.contiguous()
    out = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16) if out is None else out
    n = x.numel()
    if n % 4 != 0:
        out.copy_(x)
        return out
    check(lib.fr_cast_f32_to_bf16(x.data_ptr(), out.data_ptr(), n, stream_ptr(x.device)), "fr_cast_f32_to_bf16")
    count_launches(1)
    return out


# ---- user encoder / loss / gradient GEMMs (CUDA only) -------------------------------------------------------
def timestamp(ring: torch.Tensor, ctr: torch.Tensor) -> None:
    check(require_native().fr_timestamp(ring.data_ptr(), ring.numel(), ctr.data_ptr(), stream_ptr(ring.device)), "fr_timestamp")
    count_launches(1)


def hostio_fetch(dst: torch.Tensor, ring_dev: int, n: int, slots: int, ctr: torch.Tensor) -> None:
    """``dst[:n] = ring[*ctr % slots]; *ctr += 1`` with ``ring`` in host-mapped pinned memory (engine/fused.py run_host)."""
    check(require_native().fr_hostio_fetch(dst.data_ptr(), ring_dev, n, slots, ctr.data_ptr(), stream_ptr(dst.device)),
          "fr_hostio_fetch")
    count_launches(1)


def hostio_publish(out_dev: int, value: torch.Tensor, slots: int, ctr: torch.Tensor) -> None:
    check(require_native().fr_hostio_publish(out_dev, value.data_ptr(), slots, ctr.data_ptr(), stream_ptr(value.device)),
          "fr_hostio_publish")
    count_launches(1)


def counter_advance(ctr: torch.Tensor, inc: int = 1) -> None:
    lib = require_native()
    check(lib.fr_counter_advance(ctr.data_ptr(), int(inc), stream_ptr(ctr.device)), "fr_counter_advance")
    count_launches(1)


def ue_gather_dropout(news_vecs, inv_his, X, p, seed, seed_dev=None):
    lib = require_native()
    check(lib.fr_ue_gather_dropout(news_vecs.data_ptr(), inv_his.data_ptr(), X.data_ptr(), X.shape[0], X.shape[1],
                                   float(p), int(seed) & (2 ** 64 - 1), _p(seed_dev), stream_ptr(X.device)),
          "fr_ue_gather_dropout")
    count_launches(1)
    return X


def ue_scatter_dx(dX, inv_his, d_news, p, seed, noise_std=0.0, noise_seed=0, seed_dev=None):
    lib = require_native()
    check(lib.fr_ue_scatter_dx(dX.data_ptr(), inv_his.data_ptr(), d_news.data_ptr(), dX.shape[0], dX.shape[1], float(p),
                               int(seed) & (2 ** 64 - 1), float(noise_std), int(noise_seed) & (2 ** 64 - 1),
                               _p(seed_dev), stream_ptr(dX.device)), "fr_ue_scatter_dx")
    count_launches(1)
    return d_news


def ue_mha_fwd(qkv, his_ids, ctx, stats, B, L, H, dk, use_mask=False, impl: Optional[str] = None):
    """User-encoder self-attention forward. ``impl``: "tc" = tcgen05 kernel (default; ``FEDREC_UE_ATTN`` overrides),
    "cuda" = the CUDA-core kernel kept as its cross-check. Same outputs and saved statistics either way."""
    lib = require_native()
    impl = impl or os.environ.get("FEDREC_UE_ATTN", "tc")
    fn, name = (lib.fr_ue_attention_tc, "fr_ue_attention_tc") if impl == "tc" else (lib.fr_ue_mha_fwd, "fr_ue_mha_fwd")
    check(fn(qkv.data_ptr(), _p(his_ids), ctx.data_ptr(), stats.data_ptr(), B, L, H, dk, int(use_mask),
             stream_ptr(qkv.device)), name)
    count_launches(1)
    return ctx


def ue_mha_bwd(qkv, dctx, stats, his_ids, dqkv, B, L, H, dk, use_mask=False, impl: Optional[str] = None):
    """User-encoder self-attention backward; ``impl`` as in ``ue_mha_fwd`` ("tc" default, "cuda" cross-check)."""
    lib = require_native()
    impl = impl or os.environ.get("FEDREC_UE_ATTN", "tc")
    fn, name = (lib.fr_ue_attention_bwd_tc, "fr_ue_attention_bwd_tc") if impl == "tc" else (lib.fr_ue_mha_bwd, "fr_ue_mha_bwd")
    check(fn(qkv.data_ptr(), dctx.data_ptr(), stats.data_ptr(), _p(his_ids), dqkv.data_ptr(), B, L, H, dk,
             int(use_mask), stream_ptr(qkv.device)), name)
    count_launches(1)
    return dqkv


def ue_pool_score(ctx, e, w2, b2, news_vecs, inv_cand, his_ids, score, loss, user_vec, d_ctx_pool, d_pre, d_w2, d_b2,
                  d_news, B, L, C, use_mask=False, train=True, noise_std=0.0, noise_seed=0, seed_dev=None):
    lib = require_native()
    check(lib.fr_ue_pool_score(ctx.data_ptr(), e.data_ptr(), w2.data_ptr(), b2.data_ptr(), news_vecs.data_ptr(),
                               inv_cand.data_ptr(), _p(his_ids), _p(score), _p(loss), _p(user_vec), _p(d_ctx_pool),
                               _p(d_pre), _p(d_w2), _p(d_b2), _p(d_news), B, L, C, ctx.shape[1], e.shape[1],
                               int(use_mask), int(train), float(noise_std), int(noise_seed) & (2 ** 64 - 1),
                               _p(seed_dev), stream_ptr(ctx.device)), "fr_ue_pool_score")
    count_launches(1)


def colsum(X, out, m_dev=None, m_scale=1):
    """``out[n] += sum_m X[m, n]`` (X bf16 or fp32, unit inner stride)."""
    lib = require_native()
    assert X.stride(1) == 1 and out.dtype == torch.float32
    check(lib.fr_colsum(X.data_ptr(), 1 if X.dtype == torch.bfloat16 else 0, out.data_ptr(), X.shape[0], X.shape[1],
                        X.stride(0), _p(m_dev), m_scale, stream_ptr(X.device)), "fr_colsum")
    count_launches(1)
    return out


def colsum_weighted(X, row_w, out, m_dev=None, m_scale=1):
    """``out[n] += sum_m row_w[m] * X[m, n]`` (X bf16 or fp32, row_w fp32)."""
    lib = require_native()
    assert X.stride(1) == 1 and out.dtype == torch.float32 and row_w.dtype == torch.float32
    check(lib.fr_colsum_weighted(X.data_ptr(), 1 if X.dtype == torch.bfloat16 else 0, row_w.data_ptr(), out.data_ptr(),
                                 X.shape[0], X.shape[1], X.stride(0), _p(m_dev), m_scale, stream_ptr(X.device)),
          "fr_colsum_weighted")
    count_launches(1)
    return out


def transpose_cast(w: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out[c, r] = bf16(w[r, c])`` for a contiguous fp32 matrix."""
    lib = require_native()
    R, Cc = w.shape
    assert w.is_contiguous() and w.dtype == torch.float32
    out = torch.empty(Cc, R, device=w.device, dtype=torch.bfloat16) if out is None else out
    check(lib.fr_transpose_cast(w.data_ptr(), out.data_ptr(), R, Cc, out.stride(0), stream_ptr(w.device)), "fr_transpose_cast")
    count_launches(1)
    return out


def gemm_grad(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, k_dev=None, k_scale=1, splits=0) -> torch.Tensor:
    """``out[M1, N1] += A[K, M1]^T @ B[K, N1]`` on tcgen05 (MN-major operands, split-K, fp32 accumulate into ``out``)."""
    lib = require_native()
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and out.dtype == torch.float32
    assert A.stride(1) == 1 and B.stride(1) == 1 and out.stride(1) == 1 and A.shape[0] == B.shape[0]
    check(lib.fr_gemm_grad_bf16(A.data_ptr(), B.data_ptr(), out.data_ptr(), A.shape[1], B.shape[1], A.shape[0],
                                A.stride(0), B.stride(0), out.stride(0), _p(k_dev), k_scale, splits,
                                stream_ptr(A.device)), "fr_gemm_grad_bf16")
    count_launches(1)
    return out


def text_pool_bwd(hidden, e, alpha, d_pooled, w2, count, n_items, T, d_pre, da_scratch, d_w2, d_b2, d_b1):
    """``d_w2`` / ``d_b1`` may be None: run ``colsum_weighted(e, da_scratch, d_w2)`` and ``colsum(d_pre, d_b1)`` yourself."""
    lib = require_native()
    check(lib.fr_text_pool_bwd(hidden.data_ptr(), e.data_ptr(), alpha.data_ptr(), d_pooled.data_ptr(), w2.data_ptr(),
                               _p(count), n_items, T, e.shape[1], hidden.shape[1], d_pre.data_ptr(),
                               da_scratch.data_ptr(), _p(d_w2), d_b2.data_ptr(), _p(d_b1),
                               stream_ptr(hidden.device)), "fr_text_pool_bwd")
    count_launches(1 + (d_w2 is not None) + (d_b1 is not None))
    return d_pre


def cache_select(uniq, count, n_items, valid, miss, miss_count):
    lib = require_native()
    check(lib.fr_cache_select(uniq.data_ptr(), _p(count), n_items, valid.data_ptr(), miss.data_ptr(), miss_count.data_ptr(),
                              stream_ptr(uniq.device)), "fr_cache_select")
    count_launches(1)


def rows_copy(src, src_idx, dst, dst_idx, count, n_items, rows_per_item, valid=None):
    """Item-block row copy with optional gather (``src_idx``) / scatter (``dst_idx``); 2-D tensors of equal row width."""
    lib = require_native()
    assert src.shape[1] == dst.shape[1] and src.dtype == dst.dtype and src.stride(1) == 1 and dst.stride(1) == 1
    check(lib.fr_rows_copy(src.data_ptr(), _p(src_idx), dst.data_ptr(), _p(dst_idx), _p(count), n_items, rows_per_item,
                           src.shape[1] * src.element_size(), _p(valid), stream_ptr(src.device)), "fr_rows_copy")
    count_launches(1)
