"""fedrec_b200: a Blackwell (sm_100a) native federated news-recommendation framework.

Capabilities follow VishnoiAman777/FedRec-with-PytorchDistributed;
the architecture does not: one process per GPU, a flat fp32 master buffer for the
trainable state living in peer-mapped memory, hand-written sm_100a kernels for the
hot ops, and aggregation kernels that move data over NVLink from inside the kernel.
"""
__version__ = "0.1.0"

from .config import FedRecConfig  # noqa: F401
