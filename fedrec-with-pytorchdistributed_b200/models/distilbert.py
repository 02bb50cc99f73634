"""A self-contained DistilBERT trunk with HuggingFace-compatible parameter names.

The reference wraps ``transformers.DistilBertModel`` (encoder.py:19,27).  This
framework owns the trunk instead, for three reasons: the weights must be packed in
bf16 for the tcgen05 GEMMs, the forward must run on the framework's own kernels
(``ops.bert``), and there is no network to download a checkpoint from, so the
module has to be constructible from a config.  Key names and shapes match
``DistilBertModel.state_dict()`` exactly, so a reference
checkpoint loads unchanged; ``tests/test_distilbert.py`` checks the fp32 forward
against the installed ``transformers`` implementation.

Architecture (DistilBERT-base): word + position embeddings → LayerNorm(1e-12) →
6 × [MHSA(12×64, q scaled by 1/8, additive −inf key mask, softmax) → +res → LN →
Linear 768→3072 → GELU(erf) → Linear 3072→768 → +res → LN].
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
from torch import nn
import torch.nn.functional as F


@dataclass
class DistilBertConfigLite:
    vocab_size: int = 30522
    max_position_embeddings: int = 512
    n_layers: int = 6
    n_heads: int = 12
    dim: int = 768
    hidden_dim: int = 3072
    dropout: float = 0.1
    attention_dropout: float = 0.1
    hidden_size: int = 768       # alias used by encoder.py:20 (`config.hidden_size`)
    pad_token_id: int = 0
    initializer_range: float = 0.02


class _Embeddings(nn.Module):
    def __init__(self, c: DistilBertConfigLite):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.dim, padding_idx=c.pad_token_id)
        self.position_embeddings = nn.Embedding(c.max_position_embeddings, c.dim)
        self.LayerNorm = nn.LayerNorm(c.dim, eps=1e-12)
        self.dropout = nn.Dropout(c.dropout)

    def forward(self, input_ids: torch.Tensor) -> torch.Tensor:
        T = input_ids.shape[1]
        pos = torch.arange(T, device=input_ids.device)
        x = self.word_embeddings(input_ids) + self.position_embeddings(pos)[None]
        return self.dropout(self.LayerNorm(x))


class _SelfAttention(nn.Module):
    def __init__(self, c: DistilBertConfigLite):
        super().__init__()
        self.n_heads, self.dim = c.n_heads, c.dim
        self.q_lin = nn.Linear(c.dim, c.dim)
        self.k_lin = nn.Linear(c.dim, c.dim)
        self.v_lin = nn.Linear(c.dim, c.dim)
        self.out_lin = nn.Linear(c.dim, c.dim)
        self.dropout = nn.Dropout(c.attention_dropout)
        self.use_sdpa = False          # library attention (F.scaled_dot_product_attention) for the strong-baseline bench arm

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor]) -> torch.Tensor:
        n, T, D = x.shape
        H, dh = self.n_heads, D // self.n_heads

        def split(t):
            return t.view(n, T, H, dh).transpose(1, 2)

        if self.use_sdpa and not (self.training and self.dropout.p > 0):
            am = None if mask is None else (mask != 0)[:, None, None, :]
            ctx = F.scaled_dot_product_attention(split(self.q_lin(x)), split(self.k_lin(x)), split(self.v_lin(x)),
                                                 attn_mask=am)
            return self.out_lin(ctx.transpose(1, 2).reshape(n, T, D))

        q = split(self.q_lin(x)) / math.sqrt(dh)
        k = split(self.k_lin(x))
        v = split(self.v_lin(x))
        s = q @ k.transpose(-1, -2)
        if mask is not None:
            s = s.masked_fill((mask == 0)[:, None, None, :], torch.finfo(s.dtype).min)
        w = self.dropout(torch.softmax(s, dim=-1))
        ctx = (w @ v).transpose(1, 2).reshape(n, T, D)
        return self.out_lin(ctx)


class _FFN(nn.Module):
    def __init__(self, c: DistilBertConfigLite):
        super().__init__()
        self.lin1 = nn.Linear(c.dim, c.hidden_dim)
        self.lin2 = nn.Linear(c.hidden_dim, c.dim)
        self.dropout = nn.Dropout(c.dropout)

    def forward(self, x):
        return self.dropout(self.lin2(F.gelu(self.lin1(x))))


class _Block(nn.Module):
    def __init__(self, c: DistilBertConfigLite):
        super().__init__()
        self.attention = _SelfAttention(c)
        self.sa_layer_norm = nn.LayerNorm(c.dim, eps=1e-12)
        self.ffn = _FFN(c)
        self.output_layer_norm = nn.LayerNorm(c.dim, eps=1e-12)

    def forward(self, x, mask):
        x = self.sa_layer_norm(self.attention(x, mask) + x)
        return self.output_layer_norm(self.ffn(x) + x)


class _Transformer(nn.Module):
    def __init__(self, c: DistilBertConfigLite):
        super().__init__()
        self.layer = nn.ModuleList([_Block(c) for _ in range(c.n_layers)])

    def forward(self, x, mask):
        for blk in self.layer:
            x = blk(x, mask)
        return x


class DistilBertTrunk(nn.Module):
    """``forward(input_ids, attention_mask) -> (last_hidden_state,)`` like the HF model."""

    def __init__(self, config: Optional[DistilBertConfigLite] = None):
        super().__init__()
        self.config = config or DistilBertConfigLite()
        self.embeddings = _Embeddings(self.config)
        self.transformer = _Transformer(self.config)
        self.apply(self._init)

    def _init(self, m):
        r = self.config.initializer_range
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, std=r)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, std=r)
            if m.padding_idx is not None:
                with torch.no_grad():
                    m.weight[m.padding_idx].zero_()
        elif isinstance(m, nn.LayerNorm):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)

    @classmethod
    def from_pretrained(cls, name_or_path: str = "distilbert-base-uncased", **_):
        """A local HuggingFace-format directory (``config.json`` + ``model.safetensors`` or ``pytorch_model.bin``) if
        there is one, else random-init from the default config.

        There is no network here; the reference's ``from_pretrained`` download (encoder.py:19) becomes "use local
        weights when the user has them" (``FedRecConfig.bert_path``). Every trunk tensor has to be in the file with the
        right shape; heads the trunk does not have (``vocab_*`` of the masked-LM checkpoint) are ignored.
        """
        import os
        if not os.path.isdir(name_or_path):
            return cls()
        model = cls(read_hf_config(os.path.join(name_or_path, "config.json")))
        sd = None
        if os.path.isfile(os.path.join(name_or_path, "model.safetensors")):
            sd = read_safetensors(os.path.join(name_or_path, "model.safetensors"))
        else:
            for fname in ("pytorch_model.bin", "model.pt"):
                if os.path.isfile(os.path.join(name_or_path, fname)):
                    sd = torch.load(os.path.join(name_or_path, fname), map_location="cpu", weights_only=True)
                    break
        if sd is None:
            raise FileNotFoundError(f"{name_or_path}: no model.safetensors, pytorch_model.bin or model.pt")
        sd = {k.removeprefix("distilbert."): v for k, v in sd.items()}
        want = model.state_dict()
        missing = [k for k in want if k not in sd]
        wrong = [k for k in want if k in sd and tuple(sd[k].shape) != tuple(want[k].shape)]
        if missing or wrong:
            raise ValueError(f"{name_or_path}: {len(missing)} trunk tensors missing (e.g. {missing[:3]}), {len(wrong)} with "
                             f"a different shape (e.g. {wrong[:3]}) - is config.json the one of these weights?")
        model.load_state_dict({k: sd[k].to(want[k].dtype) for k in want})
        model.loaded_from = name_or_path
        return model

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None):
        x = self.embeddings(input_ids)
        return (self.transformer(x, attention_mask),)


def read_hf_config(path: str) -> DistilBertConfigLite:
    """``config.json`` of a DistilBERT checkpoint → the fields this trunk uses (defaults = distilbert-base-uncased)."""
    import json
    import os
    c = DistilBertConfigLite()
    if not os.path.isfile(path):
        return c
    with open(path, encoding="utf-8") as f:
        j = json.load(f)
    mt = j.get("model_type", "distilbert")
    if mt != "distilbert":
        raise ValueError(f"{path}: model_type {mt!r}; this trunk implements DistilBERT")
    for name in ("vocab_size", "max_position_embeddings", "n_layers", "n_heads", "dim", "hidden_dim", "dropout",
                 "attention_dropout", "pad_token_id", "initializer_range"):
        if name in j:
            setattr(c, name, type(getattr(c, name))(j[name]))
    if j.get("activation", "gelu") != "gelu" or j.get("sinusoidal_pos_embds", False):
        raise ValueError(f"{path}: only the GELU / learned-position DistilBERT variant is implemented")
    c.hidden_size = c.dim
    return c


_ST_DTYPES = {"F64": torch.float64, "F32": torch.float32, "F16": torch.float16, "BF16": torch.bfloat16,
              "I64": torch.int64, "I32": torch.int32, "I16": torch.int16, "I8": torch.int8, "U8": torch.uint8,
              "BOOL": torch.bool}


def read_safetensors(path: str) -> dict:
    """The safetensors container: u64 little-endian header length, a JSON header ``name → {dtype, shape, data_offsets}``,
    then the raw little-endian tensor bytes. Read with numpy's memmap, so only what is copied out is paged in."""
    import json
    import numpy as np
    with open(path, "rb") as f:
        n = int.from_bytes(f.read(8), "little")
        if n <= 0 or n > 100 << 20:
            raise ValueError(f"{path}: not a safetensors file (header length {n})")
        header = json.loads(f.read(n).decode("utf-8"))
    raw = np.memmap(path, dtype=np.uint8, mode="r", offset=8 + n)
    out = {}
    for name, meta in header.items():
        if name == "__metadata__":
            continue
        dt = _ST_DTYPES.get(meta["dtype"])
        if dt is None:
            raise ValueError(f"{path}: tensor {name} has unsupported dtype {meta['dtype']}")
        lo, hi = meta["data_offsets"]
        shape = tuple(meta["shape"])
        t = torch.frombuffer(bytearray(raw[lo:hi]), dtype=dt) if hi > lo else torch.empty(0, dtype=dt)
        out[name] = t.reshape(shape)
    return out
