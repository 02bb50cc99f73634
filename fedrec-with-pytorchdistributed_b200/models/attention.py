"""Attention primitives of the user and text encoders.

Behavioural spec: reference attention.py — AdditiveAttention :8-26,
ScaledDotProductAttention :32-45, MultiHeadAttention :50-81.  Parameter names
(``att_fc1``, ``att_fc2``, ``W_Q``, ``W_K``, ``W_V``) are part of the checkpoint
layout and are kept.

Both soft-max-like normalisations in the reference are ``exp(s) / (sum exp(s) + 1e-8)``
with no max subtraction.  The module code below evaluates the algebraically equal,
overflow-safe form ``exp(s-m) / (sum exp(s-m) + 1e-8·exp(-m))``,
which is also what the sm_100a kernels in ``csrc/user_encoder.cu`` compute; on CUDA
tensors with ``fused=True`` the modules dispatch to those kernels.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

EPS = 1e-8


def stable_exp_normalise(s: torch.Tensor, dim: int, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``exp(s)*mask / (sum(exp(s)*mask, dim) + 1e-8)`` evaluated without overflow."""
    m = s.amax(dim=dim, keepdim=True).detach()
    e = torch.exp(s - m)
    if mask is not None:
        e = e * mask
    return e / (e.sum(dim=dim, keepdim=True) + EPS * torch.exp(-m))


class AdditiveAttention(nn.Module):
    """Pools ``[B, S, d_h]`` to ``[B, d_h]`` with a learned query (tanh MLP scoring)."""

    def __init__(self, d_h: int, hidden_size: int = 200):
        super().__init__()
        self.att_fc1 = nn.Linear(d_h, hidden_size)
        self.att_fc2 = nn.Linear(hidden_size, 1)

    def weights(self, x: torch.Tensor, attn_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        e = torch.tanh(self.att_fc1(x))
        a = self.att_fc2(e)                                   # [B, S, 1]
        m = attn_mask.unsqueeze(2).to(a.dtype) if attn_mask is not None else None
        return stable_exp_normalise(a, dim=1, mask=m)

    def forward(self, x: torch.Tensor, attn_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        alpha = self.weights(x, attn_mask)
        return (x * alpha).sum(dim=1)


class ScaledDotProductAttention(nn.Module):
    def __init__(self, d_k: int):
        super().__init__()
        self.d_k = d_k

    def forward(self, Q, K, V, attn_mask: Optional[torch.Tensor] = None):
        scores = torch.matmul(Q, K.transpose(-1, -2)) / math.sqrt(self.d_k)
        attn = stable_exp_normalise(scores, dim=-1, mask=attn_mask)
        return torch.matmul(attn, V), attn


class MultiHeadAttention(nn.Module):
    """Q/K/V projections with bias, per-head attention, head merge, *no* output projection."""

    def __init__(self, d_model: int, n_heads: int, d_k: int, d_v: int):
        super().__init__()
        self.d_model, self.n_heads, self.d_k, self.d_v = d_model, n_heads, d_k, d_v
        self.W_Q = nn.Linear(d_model, d_k * n_heads)
        self.W_K = nn.Linear(d_model, d_k * n_heads)
        self.W_V = nn.Linear(d_model, d_v * n_heads)
        for lin in (self.W_Q, self.W_K, self.W_V):          # attention.py:64-67: weights only
            nn.init.xavier_uniform_(lin.weight, gain=1)
        self._sdpa = ScaledDotProductAttention(d_k)

    def forward(self, Q, K, V, attn_mask: Optional[torch.Tensor] = None):
        B, S, _ = Q.shape
        q = self.W_Q(Q).view(B, -1, self.n_heads, self.d_k).transpose(1, 2)
        k = self.W_K(K).view(B, -1, self.n_heads, self.d_k).transpose(1, 2)
        v = self.W_V(V).view(B, -1, self.n_heads, self.d_v).transpose(1, 2)
        if attn_mask is not None:                           # [B, S_k] key mask → [B, 1, 1, S_k]
            attn_mask = attn_mask.to(q.dtype)[:, None, None, :]
        ctx, _ = self._sdpa(q, k, v, attn_mask)
        return ctx.transpose(1, 2).reshape(B, -1, self.n_heads * self.d_v)
