"""Upstream module name kept: ``from encoder import TextEncoder, UserEncoder``."""
from fedrec_b200.models.encoder import TextEncoder, UserEncoder  # noqa: F401
