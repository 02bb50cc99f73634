"""Operator layer.

Every op has (a) a plain PyTorch fp32 definition in ``ops.reference`` — the oracle
the GPU tests compare against, and the path that runs on CPU — and (b) a
hand-written sm_100a kernel in ``csrc/`` reached through the C ABI of
``libfedrec_b200.so`` (``ops.native``).  On a CUDA tensor the native kernel is
mandatory: a missing or unloadable library raises instead of silently falling
back to library kernels.
"""
from . import reference  # noqa: F401
from .native import has_native, require_native, native_lib_path  # noqa: F401
from .dispatch import (  # noqa: F401
    adam_flat, segment_sum, score_ce, clip_noise_, grad_sq_norm,
)
