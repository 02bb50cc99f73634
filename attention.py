"""Upstream module name kept: ``from attention import AdditiveAttention, ...``."""
from fedrec_b200.models.attention import (  # noqa: F401
    AdditiveAttention, ScaledDotProductAttention, MultiHeadAttention)
