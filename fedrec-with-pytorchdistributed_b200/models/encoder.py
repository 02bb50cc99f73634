"""Text and user encoders.

Behavioural spec: reference encoder.py — TextEncoder :12-30, UserEncoder :36-56.
The attribute name ``DistillBert`` (two l's) is part of every checkpoint key
 and is kept.

The text encoder is split into ``trunk`` (frozen DistilBERT, model.py:25-26) and
``head`` (trainable additive attention + 768→400 projection) because the training
engine caches trunk activations: with the trunk frozen, the reference's second
"re-forward with grad" pass (model.py:72-90) only ever needs gradients through
the head.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn
import torch.nn.functional as F

from .attention import AdditiveAttention, MultiHeadAttention
from .distilbert import DistilBertTrunk, DistilBertConfigLite


class TextEncoder(nn.Module):
    def __init__(self, word_embedding_dim: int = 400, dropout_rate: float = 0.2, enable_gpu: bool = True,
                 bert_config: Optional[DistilBertConfigLite] = None, pretrained: str = "distilbert-base-uncased",
                 mask_padding: bool = False):
        super().__init__()
        self.dropout_rate = dropout_rate          # set but never applied upstream (encoder.py:18)
        self.mask_padding = mask_padding          # reference pools padded tokens too
        self.DistillBert = DistilBertTrunk(bert_config) if bert_config is not None \
            else DistilBertTrunk.from_pretrained(pretrained)
        hid = self.DistillBert.config.hidden_size
        self.additive_attention = AdditiveAttention(hid, hid // 2)
        self.fc = nn.Linear(hid, word_embedding_dim)

    def trunk(self, tokens: torch.Tensor, atts: Optional[torch.Tensor]) -> torch.Tensor:
        return self.DistillBert(tokens, attention_mask=atts)[0]

    def head(self, hidden: torch.Tensor, atts: Optional[torch.Tensor] = None) -> torch.Tensor:
        pooled = self.additive_attention(hidden, atts if self.mask_padding else None)
        return self.fc(pooled)

    def forward(self, text: torch.Tensor) -> torch.Tensor:
        """``text`` is ``[n, 2, T]``: row 0 token ids, row 1 attention mask."""
        tokens, atts = text[:, 0, :], text[:, 1, :]
        return self.head(self.trunk(tokens, atts), atts)


class UserEncoder(nn.Module):
    def __init__(self, news_embedding_dim: int = 400, num_attention_heads: int = 20,
                 query_vector_dim: int = 200, head_dim: int = 20, dropout_rate: float = 0.2,
                 mask_padding: bool = False):
        super().__init__()
        self.dropout_rate = dropout_rate
        self.mask_padding = mask_padding
        self.multihead_attention = MultiHeadAttention(news_embedding_dim, num_attention_heads, head_dim, head_dim)
        self.additive_attention = AdditiveAttention(num_attention_heads * head_dim, query_vector_dim)

    def forward(self, clicked_news_vecs: torch.Tensor, his_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = F.dropout(clicked_news_vecs, p=self.dropout_rate, training=self.training)
        m = his_mask if self.mask_padding else None
        ctx = self.multihead_attention(x, x, x, m)
        return self.additive_attention(ctx, m)
