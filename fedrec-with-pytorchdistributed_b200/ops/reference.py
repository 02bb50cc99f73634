"""Plain PyTorch fp32 definitions of every op: the oracle for GPU tests and the CPU path."""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

EPS = 1e-8


def score_ce(cand: torch.Tensor, user: torch.Tensor, labels: Optional[torch.Tensor] = None):
    """``cand [B,C,D]``, ``user [B,D]`` → ``(loss, score[B,C])`` with sigmoid *before* the soft-max CE
    (reference model.py:121-126)."""
    score = torch.sigmoid(torch.bmm(cand, user.unsqueeze(-1)).squeeze(-1))
    if labels is None:
        labels = torch.zeros(cand.shape[0], dtype=torch.long, device=cand.device)
    return F.cross_entropy(score, labels.long()), score


def segment_sum(vals: torch.Tensor, seg: torch.Tensor, num_segments: int) -> torch.Tensor:
    out = torch.zeros(num_segments, vals.shape[1], dtype=vals.dtype, device=vals.device)
    return out.index_add_(0, seg.long(), vals)


def adam_(p, g, m, v, step: int, lr: float, b1: float, b2: float, eps: float, grad_scale: float = 1.0) -> None:
    """One in-place Adam step with ``torch.optim.Adam`` default semantics; ``step`` is the 1-based count."""
    g = g * grad_scale
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


def dedup(ids: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    return torch.unique(ids.reshape(-1), return_inverse=True)


def clip_noise_(x: torch.Tensor, clip: float, noise_std: float, generator=None) -> torch.Tensor:
    if clip > 0:
        nrm = x.norm()
        x.mul_(torch.clamp(clip / nrm.clamp_min(1e-30), max=1.0))
    if noise_std > 0:
        x.add_(torch.randn(x.shape, device=x.device, dtype=x.dtype, generator=generator) * noise_std)
    return x


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], epi: int = 0,
           resid: Optional[torch.Tensor] = None) -> torch.Tensor:
    y = x.float() @ w.float().t()
    if b is not None:
        y = y + b.float()
    if epi == 1:
        y = F.gelu(y)
    elif epi == 2:
        y = y + resid.float()
    elif epi == 3:
        y = torch.tanh(y)
    return y


def exp_normalise(s: torch.Tensor, dim: int, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The reference's ``exp(s)*mask / (sum + 1e-8)`` verbatim (attention.py:19-22, :39-42)."""
    e = torch.exp(s)
    if mask is not None:
        e = e * mask
    return e / (e.sum(dim=dim, keepdim=True) + EPS)


def user_encoder(x, wq, bq, wk, bk, wv, bv, w1, b1, w2, b2, n_heads: int,
                 key_mask: Optional[torch.Tensor] = None, drop_mask: Optional[torch.Tensor] = None):
    """``x [B,L,D]`` → user vector ``[B, H*dv]`` (encoder.py:49-56 + attention.py).

    ``drop_mask`` is the already-scaled dropout multiplier (``{0, 1/(1-p)}``) or None.
    """
    B, L, D = x.shape
    if drop_mask is not None:
        x = x * drop_mask
    dk = wq.shape[0] // n_heads

    def proj(w, b):
        return (x @ w.t() + b).view(B, L, n_heads, dk).transpose(1, 2)

    q, k, v = proj(wq, bq), proj(wk, bk), proj(wv, bv)
    s = q @ k.transpose(-1, -2) / math.sqrt(dk)
    km = key_mask[:, None, None, :].to(s.dtype) if key_mask is not None else None
    a = exp_normalise(s, -1, km)
    ctx = (a @ v).transpose(1, 2).reshape(B, L, n_heads * dk)
    e = torch.tanh(ctx @ w1.t() + b1)
    al = e @ w2.t() + b2
    pm = key_mask[:, :, None].to(s.dtype) if key_mask is not None else None
    al = exp_normalise(al, 1, pm)
    return (ctx * al).sum(1)


def additive_pool(x, w1, b1, w2, b2, mask: Optional[torch.Tensor] = None):
    e = torch.tanh(x @ w1.t() + b1)
    al = e @ w2.t() + b2
    pm = mask[:, :, None].to(al.dtype) if mask is not None else None
    al = exp_normalise(al, 1, pm)
    return (x * al).sum(1)


def fold_layernorm(w: torch.Tensor, bias: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor
                   ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Weights of ``Linear(LayerNorm(x))`` rewritten so that the normalisation can be applied *after* the matmul:

        LN(x)·Wᵀ + b  =  rstd ⊙ (x·W'ᵀ)  −  (rstd·μ) ⊙ c  +  b'        (μ, rstd = row statistics of x)

    with ``W' = W∘γ`` (returned in bf16, the GEMM operand), ``c = rowsum(W')`` taken from the *rounded* W' so that a
    constant row cancels exactly, and ``b' = b + W·β``. Used by ``engine.fused.PackedTrunk`` for ``fr_gemm_bf16_fold``."""
    wf = (w.float() * gamma.float()[None, :]).bfloat16()
    return wf, wf.float().sum(1), bias.float() + w.float() @ beta.float()


def linear_after_layernorm_folded(x: torch.Tensor, wf: torch.Tensor, c: torch.Tensor, bf: torch.Tensor,
                                  eps: float = 1e-12) -> torch.Tensor:
    """fp32 statement of what the folded GEMM epilogue computes from the un-normalised ``x`` (oracle for the kernel)."""
    xf = x.float()
    mu = xf.mean(1, keepdim=True)
    rstd = torch.rsqrt((xf * xf).mean(1, keepdim=True) - mu * mu + eps)
    return rstd * (xf @ wf.float().t()) - (rstd * mu) * c[None, :] + bf[None, :]

